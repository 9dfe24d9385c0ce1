#!/bin/bash
# profiles/build_variant.sh <name> "<extra nvcc flags>" : a second build of the library with other compile-time
# parameters, written to profiles/variants/libnbody_b200_<name>.so (git-ignored, shipped by gpurun).  Select it with
# NB_LIB=<path> (n-body-simulation_b200/__init__.py).  Tuning experiments only; the product is the default build.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="$HERE/../n-body-simulation_b200/csrc"
OUT="$HERE/variants"
mkdir -p "$OUT"
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=default -ccbin /usr/bin/g++ $2"
pids=()
for f in api direct calib bh bh3d integrate stats ics; do
  /usr/local/cuda/bin/nvcc $FLAGS -c -o "$TMP/$f.o" "$SRC/$f.cu" 2>/dev/null &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
/usr/local/cuda/bin/nvcc -shared -o "$OUT/libnbody_b200_$1.so" "$TMP"/*.o -ccbin /usr/bin/g++ -lcudart 2>/dev/null
echo "built $OUT/libnbody_b200_$1.so"
