"""Small invocation of every kernel path, sized for `compute-sanitizer --tool memcheck` (closed on this round's pool: the
script was run plain, exit 0; the parity tests are the memory check that could be run):
direct (split, ordered, fused step), Barnes-Hut 2D (plain, instrumented, depth-first fallback, fused step in place / rows /
peer pointers, ragged sizes), Barnes-Hut 3D, steppers, stats, initial conditions."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
I = pkg.integrators
eng = pkg.Engine(0, 1024)
for n in (1, 33, 2049, 5000, 20001):
    pos = torch.from_numpy(pkg.ics.uniform_disk(n, seed=n, z_half_height=100.0)).cuda()
    pos2 = torch.from_numpy(pkg.ics.clustered(n, seed=n, sigma=30.0) if n > 100 else pkg.ics.uniform_disk(n, seed=n)).cuda()
    vel = torch.zeros(n, 4, device="cuda")
    acc = torch.zeros(n, 4, device="cuda")
    for dim in (2, 3):
        eng.direct_accel(pos, acc, dim, 1.01)
        eng.direct_accel(pos, acc, dim, 1.01, pkg.NB_DIRECT_ORDERED)
        eng.direct_accel(pos, acc, dim, 1.01, pkg.NB_DIRECT_NO_HOIST)
        eng.direct_step_slice(pos, vel, dim, 1.01, 1e-3, 0, n, pre_drift=(I.DRIFT_HALF_F64, 0.0), kick=(I.KICK_F32, 0.0),
                              drift=(I.DRIFT_HALF_F64, 0.0))
    eng.bh_accel(pos2, acc, 0.5, 1.01)
    eng.bh_accel(pos2, acc, 0.5, 1.01, pkg.NB_BH_COUNT_WALK)
    os.environ["NB_BH_WALK_QCAP"] = "4"
    eng.bh_accel(pos2, acc, 0.3, 1.01, pkg.NB_BH_COUNT_WALK)
    del os.environ["NB_BH_WALK_QCAP"]
    groups = (n + 31) // 32
    kw = dict(pre_drift=(I.DRIFT_HALF_F64, 0.0), kick=(I.KICK_F32, 0.0), drift=(I.DRIFT_HALF_F64, 0.0))
    eng.bh_step_groups(pos2, vel, 0.5, 1.01, 1e-3, 0, groups, **kw)
    rows = torch.zeros(32 * groups, 4, device="cuda")
    eng.bh_step_groups(pos2, vel, 0.5, 1.01, 1e-3, 0, groups, rows_out=rows, **kw)
    eng.bh_step_groups(pos2, vel, 0.5, 1.01, 1e-3, 0, groups, rows_in=rows, prev_drift=(I.DRIFT_HALF_F64, 0.0), rows_out=rows, **kw)
    buf = eng.peer_alloc((32 * groups, 4))
    eng.bh_step_groups(pos2, vel, 0.5, 1.01, 1e-3, 0, groups, rows_in=rows, prev_drift=(I.DRIFT_HALF_F64, 0.0),
                       rows_peers=[buf.ptr, buf.ptr], flags=pkg.NB_BH_PLANAR, **kw)
    eng.bh_step_groups(pos2, vel, 0.5, 1.01, 1e-3, 0, 0, rows_in=buf.tensor, prev_drift=(I.DRIFT_HALF_F64, 0.0))
    torch.cuda.synchronize()
    buf.release()
    eng.bh3d_accel(pos, acc, 0.5, 1.01)
    eng.bh3d_accel(pos, acc, 0.5, 1.01, pkg.NB_BH_COUNT_WALK)
    eng.stats(pos, vel, 2)
    eng.ic_standard(pos, 0, n, seed=3)
    eng.ic_standard_velocities(pos, acc, vel, 0, n, seed=3)
    eng.ic_uniform_ellipsoid(pos, 0, n, seed=3)
    torch.cuda.synchronize()
print("SANITIZE_SCRIPT_DONE")
