"""Rows f-1 and f-2 of SURVEY.md section 8 pinned to THE REFERENCE ITSELF, not to a restatement: oracle/_ref/nbody_ref_prog
is the reference's unchanged NbodySystem + IntegrationMethods + SimulationStats + its own direct kernels (oracle/
build_dropin.sh).  Both sides start from the same .params checkpoint (the reference's own loader, src/NbodySystem.cpp:27-70),
run the same number of steps of the same stepper, and

  * the trajectory of the mirrored steppers (integrators.py on nb_integrate_substep, force = the direct sum in
    NB_DIRECT_ORDERED mode, which reproduces the reference kernel bit for bit at this size) must equal the reference
    program's final positions and velocities BIT FOR BIT -- every kick / drift expression type, FP64 promotions included;
  * nb_stats of that final state must equal the moment, kinetic and potential energy the reference program prints."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "nbody_ref_prog")
EPS, DT = 1.01, 0.001          # oracle/dropin_main.cu (src/main.cu:7,12)
STEPPERS = [("verlet", "Verlet_DKD"), ("euler_kd", "Euler_Symplectic_KD"), ("euler_dk", "Euler_Symplectic_DK"),
            ("vverlet", "Velocity_Verlet"), ("forestruth", "ForestRuth"), ("pefrl", "PEFRL")]


def _reference_run(tmp_path, pkg, pos, vel, stepper, steps):
    ck = str(tmp_path / "start.params")
    pkg.checkpoint.save_params(ck, pos, vel)
    out = str(tmp_path / "ref.bin")
    r = subprocess.run([REF, "direct", "0", str(steps), "0.5", out, stepper, "load:" + ck], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    m = re.search(r"STATS M0=(\S+) K0=(\S+) P0=(\S+) M1=(\S+) K1=(\S+) P1=(\S+)", r.stdout)
    assert m, r.stdout[-2000:]
    n = pos.shape[0]
    state = np.fromfile(out, dtype=np.float32).reshape(2, n, 4)
    return state[0], state[1], [float(x) for x in m.groups()]


@pytest.mark.parametrize("stepper,cls", STEPPERS)
def test_stepper_trajectory_and_stats_equal_the_reference_program(tmp_path, engine, pkg, stepper, cls):
    import torch
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/nbody_ref_prog not built (needs /root/reference at build time)")
    n, steps = 4096, 20                    # a multiple of 128: the reference kernels have no tail guard
    pos_h = pkg.ics.uniform_disk(n, seed=71)
    rng = np.random.default_rng(71)
    vel_h = np.zeros((n, 4), np.float32)
    vel_h[:, :2] = (rng.standard_normal((n, 2)) * 3.0e4).astype(np.float32)
    ref_pos, ref_vel, st = _reference_run(tmp_path, pkg, pos_h, vel_h, stepper, steps)

    pos, vel = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    m0, k0, p0 = engine.stats(pos, vel, 2)
    for got, want in zip((m0, k0, p0), st[:3]):
        assert abs(got - want) <= 1e-11 * abs(want), ("initial stats", got, want)
    sysm = pkg.integrators.ShardedSystem(engine, pos, vel, "direct", 2, EPS, direct_flags=pkg.NB_DIRECT_ORDERED)
    integ = getattr(pkg.IntegrationMethods, cls)()
    integ.setSystem(sysm)
    for _ in range(steps):
        integ.integrate(DT)
    torch.cuda.synchronize()
    got_pos, got_vel = sysm.pos.cpu().numpy(), sysm.vel.cpu().numpy()
    assert np.array_equal(got_pos[:, :2], ref_pos[:, :2]), np.abs(got_pos[:, :2] - ref_pos[:, :2]).max()
    assert np.array_equal(got_vel[:, :2], ref_vel[:, :2]), np.abs(got_vel[:, :2] - ref_vel[:, :2]).max()
    assert np.array_equal(got_pos[:, 3], ref_pos[:, 3])
    m1, k1, p1 = engine.stats(sysm.pos, sysm.vel, 2)
    for got, want in zip((m1, k1, p1), st[3:]):
        assert abs(got - want) <= 1e-11 * abs(want), ("final stats", got, want)
