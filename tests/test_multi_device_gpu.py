"""Engines on two devices in ONE process (the sharded runs use one process per GPU, which hides per-device state kept
per process: cudaFuncSetAttribute for the direct kernel's 49,200-byte shared-memory ring is per device)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_engines_on_two_devices_in_one_process(pkg, oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    n = 20000
    pos3 = pkg.ics.uniform_disk(n, seed=3, z_half_height=3000.0)
    pos2 = pkg.ics.uniform_disk(n, seed=4)
    want3 = oracle.direct_f64(pos3, 3, 1.01)
    wantb = oracle.bh_accel(pos2, 0.5, 1.01)["acc"]
    engines = [pkg.Engine(d, n) for d in (0, 1)]       # device 0 first: fills any process-wide cache
    for rep in range(2):
        for d, eng in enumerate(engines):
            dev = torch.device("cuda", d)
            with torch.cuda.device(dev):
                p3 = torch.from_numpy(pos3).to(dev)
                a3 = torch.full((n, 4), 7.0, device=dev)
                eng.direct_accel(p3, a3, 3, 1.01)
                eng.direct_accel(p3, a3, 3, 1.01, pkg.NB_DIRECT_NO_HOIST)
                p2 = torch.from_numpy(pos2).to(dev)
                a2 = torch.full((n, 4), 7.0, device=dev)
                eng.bh_accel(p2, a2, 0.5, 1.01)
                v = torch.zeros(n, 4, device=dev)
                eng.verlet_dkd_drift(p2, v, 0, n, 1e-3)
                torch.cuda.synchronize(dev)
                e3 = np.linalg.norm(a3.cpu().numpy()[:, :3] - want3[:, :3], axis=1) / np.linalg.norm(want3[:, :3], axis=1)
                eb = np.linalg.norm(a2.cpu().numpy()[:, :2] - wantb[:, :2], axis=1) / np.linalg.norm(wantb[:, :2], axis=1)
                assert e3.max() < 5e-5 and np.median(eb) < 2e-6, (d, rep, e3.max(), np.median(eb))
    for eng in engines:
        eng.close()
    # the calls on device 1 (and closing its engine) left the process's current device where they found it
    assert torch.cuda.current_device() == 0
    probe = torch.zeros(1).cuda()
    assert probe.device.index == 0
