"""GPU tests of the 3D (octree) Barnes-Hut path.  The reference has no 3D Barnes-Hut (it downgrades R3 to R2,
src/ComputeMethods.cpp:35-38), so the checker is the CPU octree of oracle/ (the restated 2D rules one dimension up)
and the FP64 direct sum with the reference's 3D force law."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = 1.01


def _cases(pkg):
    dup = pkg.ics.uniform_disk(3000, seed=91, z_half_height=3000.0)
    dup[100:140, :3] = dup[99, :3]            # coincident bodies -> depth-cap bucket
    rng = np.random.default_rng(9)
    ball = np.zeros((6000, 4), np.float32)
    ball[:, :3] = (rng.standard_normal((6000, 3)) * 5000).astype(np.float32)
    ball[:, 3] = rng.uniform(1.0, 1e5, 6000).astype(np.float32)
    return {
        "thick_disk_8192": pkg.ics.uniform_disk(8192, seed=1, z_half_height=3000.0),
        "ragged_5000": pkg.ics.uniform_disk(5000, seed=92, z_half_height=20000.0),
        "gaussian_ball_random_masses": ball,
        "duplicates_3000": dup,
        "flat_disk_in_3d": pkg.ics.uniform_disk(4097, seed=93),          # z = 0 for every body: degenerate extent
        "tiny_2": pkg.ics.uniform_disk(2, seed=94, z_half_height=10.0),
        "tiny_33": pkg.ics.uniform_disk(33, seed=95, z_half_height=3000.0),
    }


CASES = ["thick_disk_8192", "ragged_5000", "gaussian_ball_random_masses", "duplicates_3000", "flat_disk_in_3d", "tiny_2",
         "tiny_33"]


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _rel(a, b):
    return (np.linalg.norm(a[:, :3].astype(np.float64) - b[:, :3], axis=1) / np.maximum(np.linalg.norm(b[:, :3], axis=1), 1e-30))


@pytest.mark.parametrize("case", CASES)
def test_octree_stages_topology_and_accelerations(engine, pkg, oracle, case):
    import torch
    pos = _cases(pkg)[case]
    n = pos.shape[0]
    st = engine.bh3d_stages(_dev(pos))
    b = oracle.bh3_bounds(pos)
    assert np.array_equal(st["bounds"], b)
    k = oracle.bh3_tree_keys(pos, b)
    assert np.array_equal(st["tree_keys"], k)
    _, si = oracle.sort_pairs(k)
    assert np.array_equal(st["sorted_idx"], si)
    acc = torch.full((n, 4), 7.0, device="cuda")
    engine.bh3d_accel(_dev(pos), acc, 0.5, EPS, pkg.NB_BH_COUNT_WALK)
    torch.cuda.synchronize()
    got = acc.cpu().numpy()
    t = oracle.Tree3(pos, b, si)
    oc, c = t.canonical(), engine.bh3d_export_tree()
    assert np.array_equal(c["prefix"], oc["prefix"]) and np.array_equal(c["level"], oc["level"])
    assert np.array_equal(c["count"], oc["count"]) and np.allclose(c["mass"], oc["mass"], rtol=1e-12)
    scale = np.abs(pos[:, :3]).max()
    for ax in ("cx", "cy", "cz"):
        assert np.abs(c[ax] - oc[ax]).max() < 1e-9 * scale
    oa, cnt = t.walk(si, 0.5, EPS)
    assert np.all(got[:, 3] == 0)
    e = _rel(got, oa.astype(np.float64))
    assert np.median(e) < 2e-6 and np.quantile(e, 0.99) < 2e-5, (np.median(e), np.quantile(e, 0.99), e.max())
    kk = engine.bh3d_last_counters()
    assert abs(kk["interactions"] - cnt["interactions"]) <= 1e-4 * cnt["interactions"] + 64 and kk["status"] == 0


def test_octree_error_against_fp64_direct_sum(engine, pkg, oracle):
    import torch
    n = 16384
    pos = pkg.ics.uniform_disk(n, seed=96, z_half_height=6000.0)
    d = oracle.direct_f64(pos, 3, EPS)
    acc = torch.zeros(n, 4, device="cuda")
    engine.bh3d_accel(_dev(pos), acc, 1e-4, EPS)
    torch.cuda.synchronize()
    assert _rel(acc.cpu().numpy(), d).max() < 5e-5                    # theta -> 0: the direct sum
    for theta in (0.4, 0.8):
        engine.bh3d_accel(_dev(pos), acc, theta, EPS)
        torch.cuda.synchronize()
        e_gpu = _rel(acc.cpu().numpy(), d)
        e_ora = _rel(oracle.bh3_accel(pos, theta, EPS)["acc"], d)
        for q in (0.5, 0.99):
            assert abs(np.quantile(e_gpu, q) - np.quantile(e_ora, q)) <= 0.02 * np.quantile(e_ora, q) + 1e-7
        assert np.median(e_gpu) < 1e-2


def test_octree_groups_tile_and_large_n(pkg, oracle):
    import torch
    n = 1 << 20
    eng = pkg.Engine(0, n)
    pos_h = pkg.ics.uniform_disk(n, seed=1, z_half_height=3000.0)
    pos = _dev(pos_h)
    full = torch.zeros(n, 4, device="cuda")
    eng.bh3d_accel(pos, full, 0.5, EPS)
    parts = torch.full((n, 4), -1.0, device="cuda")
    groups = (n + 31) // 32
    cuts = [0, 1, 1000, groups // 2, groups]
    for a, b in zip(cuts[:-1], cuts[1:]):
        eng.bh3d_accel_groups(pos, parts, 0.5, EPS, a, b - a)
    torch.cuda.synchronize()
    assert torch.equal(full, parts)
    sample = np.arange(0, n, n // 256)[:256]
    ref = np.concatenate([oracle.direct_f64(pos_h, 3, EPS, t0=int(t), nt=1) for t in sample])
    e = _rel(full.cpu().numpy()[sample], ref)
    assert np.median(e) < 1e-2 and e.max() < 0.05, (np.median(e), e.max())   # monopole-only error of a thin 3D disc at theta/2 = 0.25
    eng.close()
