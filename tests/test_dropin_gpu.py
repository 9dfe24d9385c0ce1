"""Drop-in proof and energy drift (BASELINE.json north_star: "Total-energy drift must track the reference over
1000 steps").  oracle/build_dropin.sh links the reference's UNCHANGED NbodySystem / Verlet_DKD / DiskModel /
SimulationStats once with the reference's own ComputeMethods + kernels and once with the adapter + libnbody_b200.so;
both run the same 1000 Leapfrog steps from the reference's own initial conditions."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "nbody_ref_prog")
OURS = os.path.join(ROOT, "oracle", "_ref", "nbody_dropin_prog")


# Energy drift of the reference program measured on a B200 (1000 Leapfrog steps, DiskModel, dt = 0.001); used when
# the reference's own Barnes-Hut kernels do not finish in time: their lock-based tree build (spin on -2 with a
# __syncthreads in divergent code, src/BarnesHutKernels.cu:296-363) occasionally stalls for minutes on sm_100.
REFERENCE_DRIFT = {("direct", 4096): 2.928e-3, ("bh", 8192): 4.600e-5}


def _run(prog, method, n, steps, theta, out, integrator="verlet", ic="disk", timeout=120):
    r = subprocess.run([prog, method, str(n), str(steps), str(theta), out, integrator, ic], capture_output=True, text=True,
                       timeout=timeout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    m = re.search(r"RESULT .*E0=(\S+) E1=(\S+) drift=(\S+) moment_drift=(\S+)", r.stdout)
    assert m, r.stdout[-2000:]
    state = np.fromfile(out, dtype=np.float32).reshape(2, n, 4)
    return dict(E0=float(m.group(1)), E1=float(m.group(2)), drift=float(m.group(3)), mdrift=float(m.group(4)),
                pos=state[0], vel=state[1])


@pytest.mark.parametrize("method,n,theta", [("direct", 4096, 0.5), ("bh", 8192, 0.5)])
def test_energy_drift_tracks_reference_over_1000_steps(tmp_path, method, n, theta):
    if not (os.path.exists(REF) and os.path.exists(OURS)):
        pytest.skip("oracle/_ref drop-in programs not built (needs /root/reference at build time)")
    steps = 1000
    b = _run(OURS, method, n, steps, theta, str(tmp_path / "ours.bin"))
    try:
        a = _run(REF, method, n, steps, theta, str(tmp_path / "ref.bin"), timeout=60)
    except subprocess.TimeoutExpired:
        ref_drift = REFERENCE_DRIFT[(method, n)]
        assert abs(b["drift"] - ref_drift) <= 0.5 * abs(ref_drift) + 1e-6, (ref_drift, b["drift"])
        pytest.skip("the reference program stalled (its lock-based tree build); drift checked against the recorded value")
    # same initial state (the reference's own generator + one force evaluation for the velocities)
    assert abs(a["E0"] - b["E0"]) <= 1e-5 * abs(a["E0"])
    # drift of the same size (both are tiny: symplectic integrator, dt = 0.001)
    assert abs(b["drift"]) <= 2.0 * abs(a["drift"]) + 1e-6, (a["drift"], b["drift"])
    assert abs(b["drift"] - a["drift"]) <= 0.5 * abs(a["drift"]) + 1e-6, (a["drift"], b["drift"])
    # trajectories stay together: positions within 1e-4 of the disc radius after 1000 steps
    d = np.linalg.norm(a["pos"][:, :2] - b["pos"][:, :2], axis=1)
    assert np.quantile(d, 0.99) < 1e-4 * 30000.0, (np.median(d), d.max())
    print(f"{method}: drift ref {a['drift']:.3e} ours {b['drift']:.3e}; position diff median {np.median(d):.3e} max {d.max():.3e}")


@pytest.mark.parametrize("integrator,ic,theta", [("pefrl", "cold", 0.5), ("forestruth", "merger", 0.7)])
def test_other_integrators_and_initial_conditions_run_unchanged(tmp_path, integrator, ic, theta):
    """BASELINE.json configs 4 and 5 in miniature (N = 32,768): PEFRL on a cold collapse (4 force evaluations per
    step) and Forest-Ruth on a two-disc merger (3 per step) -- the reference's own integrator and initial-condition
    classes, once on the reference's Barnes-Hut kernels, once on this engine."""
    if not (os.path.exists(REF) and os.path.exists(OURS)):
        pytest.skip("oracle/_ref drop-in programs not built (needs /root/reference at build time)")
    n, steps = 16384, 20
    b = _run(OURS, "bh", n, steps, theta, str(tmp_path / "ours.bin"), integrator, ic)
    try:
        a = _run(REF, "bh", n, steps, theta, str(tmp_path / "ref.bin"), integrator, ic, timeout=60)
    except subprocess.TimeoutExpired:
        pytest.skip("the reference program stalled (its lock-based tree build)")
    d = np.linalg.norm(a["pos"][:, :2] - b["pos"][:, :2], axis=1)
    dv = np.linalg.norm(a["vel"][:, :2] - b["vel"][:, :2], axis=1)
    vs = np.linalg.norm(a["vel"][:, :2], axis=1)
    assert np.quantile(d, 0.999) < 1e-4 * 30000.0, (np.median(d), d.max())
    assert np.median(dv) < 1e-3 * max(np.median(vs), 1e-3), (np.median(dv), np.median(vs))
    assert abs(a["E1"] - b["E1"]) <= 1e-4 * abs(a["E1"])


def test_checkpoints_are_interchangeable_with_the_reference(tmp_path, pkg):
    """A .params file written by the reference's own CheckpointSaver is read by checkpoint.load_params, and one
    written by checkpoint.save_params is loaded by the reference's NbodySystem(configPath) constructor."""
    if not (os.path.exists(REF) and os.path.exists(OURS)):
        pytest.skip("oracle/_ref drop-in programs not built (needs /root/reference at build time)")
    n, steps = 4096, 10
    work = tmp_path / "ck"
    work.mkdir()
    out = str(tmp_path / "state.bin")
    r = subprocess.run([REF, "direct", str(n), str(steps), "0.5", out, "verlet", "disk", str(work)], capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-1000:] + r.stderr[-1000:]
    state = np.fromfile(out, dtype=np.float32).reshape(2, n, 4)
    assert open(work / "configs" / "run.json").read() == pkg.checkpoint.config_json_text(n, True)
    pos, vel = pkg.checkpoint.load_params(str(work / "configs" / f"run_{steps}.params"))
    assert np.array_equal(pos, state[0]) and np.array_equal(vel[:, :3], state[1][:, :3])
    # the other direction: our writer, the reference's loader, zero further steps -> the same state comes back
    mine = str(tmp_path / "mine.params")
    pkg.checkpoint.save_params(mine, pos, vel)
    out2 = str(tmp_path / "state2.bin")
    r = subprocess.run([OURS, "direct", "0", "0", "0.5", out2, "verlet", "load:" + mine], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stdout[-1000:] + r.stderr[-1000:]
    back = np.fromfile(out2, dtype=np.float32).reshape(2, n, 4)
    assert np.array_equal(back[0], pos) and np.array_equal(back[1][:, :3], vel[:, :3])
