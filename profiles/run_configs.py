"""The five BASELINE.json configs on ONE GPU (configs 2 and 5 are additionally sharded by bench.py / the tests):
step rates of this engine and, where the reference's kernels can run the case, of the reference's CUDA kernels on the
same box (oracle/_ref/libnbody_ref.so).  Prints one JSON line per config."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
EPS, DT = 1.01, 0.001
ref = None
so = os.path.join(ROOT, "oracle", "_ref", "libnbody_ref.so")
if os.path.exists(so):
    ref = C.CDLL(so)
    vp = C.c_void_p
    ref.ref_direct.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float]
    ref.ref_barnes_hut.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, C.c_float]


def timed_steps(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps


def ours(n, pos_h, method, integ_name, dim, theta, steps, fused=True):
    """fused: the kick / drift kernels of the stepper ride on the force kernels (nb_direct_step_slice, nb_bh_step_groups);
    else the separate kernels of round 1 (ShardedSystem + the mirrored IntegrationMethods classes)."""
    eng = pkg.Engine(0, n)
    pos = torch.from_numpy(pos_h).cuda()
    vel = torch.zeros(n, 4, device="cuda")
    if fused:
        s = (pkg.integrators.FusedDirect(eng, pos, vel, integ_name, dim, EPS) if method == "direct"
             else pkg.integrators.FusedBarnesHut(eng, pos, vel, integ_name, theta, EPS))
        dt = timed_steps(lambda: s.step(DT), steps)
        evals = len(pkg.integrators.stepper_program(integ_name, DT)[1])
    else:
        s = pkg.integrators.ShardedSystem(eng, pos, vel, method, dim, EPS, theta)
        integ = getattr(pkg.IntegrationMethods, integ_name)()
        integ.setSystem(s)
        dt = timed_steps(lambda: integ.integrate(DT), steps)
        evals = integ.force_evaluations
    eng.close()
    return dt, evals


def reference_force(n, pos_h, method, dim, theta, reps):
    """One force evaluation of the reference's kernels (the integrator kernels around it are negligible)."""
    if ref is None or n % 8192 != 0:
        return None
    pos_d = torch.zeros(3 * n, 4, device="cuda")
    pos_d[:n] = torch.from_numpy(pos_h).cuda()
    acc = torch.zeros(n, 4, device="cuda")
    if method == "direct":
        f = lambda: ref.ref_direct(pos_d.data_ptr(), acc.data_ptr(), n, dim, EPS)
    else:
        fmax = float(np.finfo(np.float32).max)
        b = torch.tensor([fmax, fmax, -fmax, -fmax], device="cuda")
        idx = torch.zeros(1, dtype=torch.int32, device="cuda")
        nodes = torch.zeros(12 * n, dtype=torch.int32, device="cuda")
        si = torch.zeros(n, dtype=torch.int32, device="cuda")
        ks = torch.zeros(n, dtype=torch.int32, device="cuda")
        f = lambda: ref.ref_barnes_hut(pos_d.data_ptr(), acc.data_ptr(), b.data_ptr(), idx.data_ptr(), nodes.data_ptr(),
                                       si.data_ptr(), ks.data_ptr(), n, 3 * n, theta, EPS)
    return timed_steps(f, reps, 1)


configs = [
    # the one timing the reference publishes (src/main.cu:11-19,53): 19.3 ms per tick on an unnamed GPU, with file output
    ("0 main.cu workload: Barnes-Hut N=32768 theta=0.04, symplectic Euler DK", 32768, "disk", "bh", "Euler_Symplectic_DK", 2, 0.04, 100),
    ("1 direct N=16384 uniform disk, Leapfrog", 16384, "disk", "direct", "Verlet_DKD", 2, 0.5, 100),
    ("2 direct N=1048576, Leapfrog (1 GPU)", 1 << 20, "disk", "direct", "Verlet_DKD", 2, 0.5, 3),
    ("3 Barnes-Hut N=1048576 theta=0.5, Leapfrog", 1 << 20, "disk", "bh", "Verlet_DKD", 2, 0.5, 20),
    ("4 Barnes-Hut cold collapse N=1048576 theta=0.5, PEFRL (late state: knots over a diffuse disc, at rest)", 1 << 20, "collapse", "bh", "PEFRL", 2, 0.5, 10),
    ("5 Barnes-Hut two-galaxy N=16777216 theta=0.7, Forest-Ruth (1 GPU)", 1 << 24, "merger", "bh", "ForestRuth", 2, 0.7, 3),
]
only = sys.argv[1:] if len(sys.argv) > 1 else None
for name, n, ic, method, integ, dim, theta, steps in configs:
    if only and name[0] not in only:
        continue
    pos_h = {"disk": lambda: pkg.ics.uniform_disk(n, seed=1), "merger": lambda: pkg.ics.two_galaxies(n),
             "collapse": lambda: pkg.ics.cold_collapse_late(n, seed=2)}[ic]()
    dt, evals = ours(n, pos_h, method, integ, dim, theta, steps)
    line = {"config": name, "ours_ms_per_step": dt * 1e3, "ours_steps_per_s": 1 / dt, "force_evaluations_per_step": evals}
    if n <= (1 << 20):
        dts, _ = ours(n, pos_h, method, integ, dim, theta, steps, fused=False)
        line["ours_ms_per_step_separate_kernels"] = dts * 1e3
    if method == "direct":
        line["ours_interactions_per_s"] = evals * float(n) * n / dt
    r = reference_force(n, pos_h, method, dim, theta, 2 if n >= (1 << 20) else 10)
    if r is not None:
        line["reference_cuda_ms_per_force_evaluation"] = r * 1e3
        line["reference_cuda_steps_per_s_estimate"] = 1 / (r * evals)
        line["speedup_vs_reference_cuda"] = (r * evals) / dt
    print(json.dumps(line), flush=True)
