"""A/B of library builds (profiles/build_variant.sh) on one GPU: each variant in its own process (NB_LIB), same inputs.
   python profiles/variants_ab.py [name ...]      (default: every .so under profiles/variants plus the product build)"""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def worker():
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    pkg = g.load_package()
    n = 1 << 20
    eng = pkg.Engine(0, n)
    out = {"lib": os.path.basename(pkg.library_path())}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timed(fn, reps, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for a, b in ev:
            flush.fill_(1)
            a.record()
            fn()
            b.record()
        torch.cuda.synchronize()
        return float(np.median([a.elapsed_time(b) for a, b in ev]))
    pos3 = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1, z_half_height=3000.0)).cuda()
    acc = torch.zeros(n, 4, device="cuda")
    peak = 148 * 128 * 2 * 1965e6 / 1e12
    only = os.environ.get("AB_ONLY", "")
    if only != "bh":
        for name, flags in (("default", 0), ("general", pkg.NB_DIRECT_NO_HOIST)):
            ms = timed(lambda: eng.direct_accel(pos3, acc, 3, 1.01, flags), 3, 1)
            out[f"direct3d_{name}_ms"] = ms
            out[f"direct3d_{name}_frac"] = float(n) * n * 20 / (ms * 1e-3) / 1e12 / peak
    pos2 = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1)).cuda()
    if only != "bh":
        out["direct2d_general_ms"] = timed(lambda: eng.direct_accel(pos2, acc, 2, 1.01, pkg.NB_DIRECT_NO_HOIST), 3, 1)
    if only != "direct":
        out["bh_eval_ms"] = timed(lambda: eng.bh_accel(pos2, acc, 0.5, 1.01), 20, 3)
        cc = pkg.ics.cold_collapse_late(n, seed=2)
        posc = torch.from_numpy(cc).cuda()
        out["bh_eval_cold_collapse_ms"] = timed(lambda: eng.bh_accel(posc, acc, 0.5, 1.01), 20, 3)
    print("AB " + json.dumps(out), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--worker":
        worker()
        sys.exit(0)
    libs = [None] + sorted(glob.glob(os.path.join(ROOT, "profiles", "variants", "*.so")))
    if len(sys.argv) > 1:
        libs = [None] + [l for l in libs[1:] if any(a in l for a in sys.argv[1:])]
    for lib in libs:
        env = dict(os.environ)
        if lib:
            env["NB_LIB"] = lib
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker"], env=env, capture_output=True, text=True)
        print((r.stdout.strip().splitlines() or ["(no output) " + r.stderr[-300:]])[-1], flush=True)
