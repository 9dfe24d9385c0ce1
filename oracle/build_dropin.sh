#!/bin/bash
# Builds the drop-in demonstration pair from the reference's sources WHERE THEY LIE (/root/reference/src), staged
# through a throw-away directory for the four MSVC->GCC portability edits of SURVEY.md section 8c.  Only binaries
# are written, into the git-ignored oracle/_ref/.  No reference source enters the repository.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REPO="$(dirname "$HERE")"
REF="${REF:-/root/reference/src}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
[ -d "$REF" ] || { echo "build_dropin: $REF absent, keeping prebuilt binaries"; exit 0; }
STAGE="$(mktemp -d)"
trap 'rm -rf "$STAGE"' EXIT
cp "$REF"/*.h "$REF"/*.cuh "$REF"/*.cpp "$REF"/*.cu "$STAGE"/
cd "$STAGE"
sed -i 's/^static struct dataBuffer/struct dataBuffer/; s/^enum Space {/enum Space : int {/' NbodySystem.h
sed -i 's/^enum Space;/enum Space : int;/' SimulationStats.cuh
sed -i '/template <>/,+1d' Callbacks.h
sed -i 's/chrono::high_resolution_clock::now()/chrono::steady_clock::now()/g' NbodySystem.cpp
rm -f main.cu
mkdir -p "$HERE/_ref"
FLAGS="-std=c++17 -O3 -gencode arch=compute_100,code=sm_100 -w -ccbin /usr/bin/g++ -x cu -I$STAGE"
COMMON="NbodySystem.cpp IntegrationMethods.cu InitialConditions.cu SimulationStats.cu Callbacks.cpp $HERE/dropin_main.cu"
# (1) the reference program: its own ComputeMethods and kernels
$NVCC $FLAGS -o "$HERE/_ref/nbody_ref_prog" $COMMON ComputeMethods.cpp DirectComputeKernels.cu BarnesHutKernels.cu
# (2) the same callers on the B200 engine: adapter + C ABI
cp "$REPO/n-body-simulation_b200/adapter/ComputeMethods.h" "$STAGE/ComputeMethods.h"
$NVCC $FLAGS -I"$REPO/include" -o "$HERE/_ref/nbody_dropin_prog" $COMMON \
      "$REPO/n-body-simulation_b200/adapter/ComputeMethods.cpp" \
      -L"$REPO/n-body-simulation_b200" -lnbody_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../../n-body-simulation_b200'
echo "build_dropin: wrote $HERE/_ref/nbody_ref_prog and nbody_dropin_prog"
