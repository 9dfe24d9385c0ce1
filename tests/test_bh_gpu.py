"""GPU parity tests of the Barnes-Hut pipeline, stage by stage, through the C ABI."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = 1.01
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = ["disk_8192", "disk_5000_ragged", "clustered_8192", "tight_clusters_4097", "duplicates_3000",
         "two_galaxies_6000", "tiny_2", "tiny_3", "tiny_33",
         "all_coincident_700", "collinear_1500", "massless_half_2048", "far_offset_4096", "all_at_origin_100", "small_knot_300", "dense_knot_6000"]


def _cases(pkg):
    dup = pkg.ics.uniform_disk(3000, seed=31)
    dup[100:140, :2] = dup[99, :2]            # 41 coincident bodies -> depth-cap bucket
    dup[2000:2003, :2] = dup[1999, :2]
    same = pkg.ics.uniform_disk(700, seed=36)
    same[:, :2] = same[1, :2]                 # tiny box (padding only): one bucket holds everything
    line = pkg.ics.uniform_disk(1500, seed=37)
    line[:, 1] = 12.5                         # zero-height box
    light = pkg.ics.uniform_disk(2048, seed=38)
    light[::2, 3] = 0.0                       # massless tracers next to massive bodies
    far = pkg.ics.uniform_disk(4096, seed=39)
    far[:, 0] += 3.0e6                        # ~0.25 resolution in x: many bodies share an x coordinate
    far[:, 1] -= 7.0e5
    rng = np.random.default_rng(40)
    knot = pkg.ics.uniform_disk(9000, seed=41)          # 300 bodies inside one level-20 cell: the sort's fix-up block
    knot[5000:5300, :2] = knot[4999, :2] + rng.uniform(-4e-3, 4e-3, (300, 2)).astype(np.float32)
    dense = pkg.ics.uniform_disk(20000, seed=42)        # 6000 of them: beyond the fix-up, the conditional full sort
    dense[7000:13000, :2] = dense[6999, :2] + rng.uniform(-4e-3, 4e-3, (6000, 2)).astype(np.float32)
    return {
        "small_knot_300": knot, "dense_knot_6000": dense,
        "all_at_origin_100": np.tile(np.array([[0.0, 0.0, 0.0, 5.0]], np.float32), (100, 1)),   # zero-extent box
        "all_coincident_700": same, "collinear_1500": line, "massless_half_2048": light, "far_offset_4096": far,
        "disk_8192": pkg.ics.uniform_disk(8192, seed=1),
        "disk_5000_ragged": pkg.ics.uniform_disk(5000, seed=32),
        "clustered_8192": pkg.ics.clustered(8192, seed=5, sigma=300.0),
        "tight_clusters_4097": pkg.ics.clustered(4097, seed=6, sigma=0.5, clusters=3),
        "duplicates_3000": dup,
        "two_galaxies_6000": pkg.ics.two_galaxies(6000),
        "tiny_2": pkg.ics.uniform_disk(2, seed=33),
        "tiny_3": pkg.ics.uniform_disk(3, seed=34),
        "tiny_33": pkg.ics.uniform_disk(33, seed=35),
    }


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _rel(a, b):
    return (np.linalg.norm(a[:, :2].astype(np.float64) - b[:, :2], axis=1)
            / np.maximum(np.linalg.norm(b[:, :2], axis=1), 1e-30))


@pytest.mark.parametrize("case", CASES)
def test_integer_stages_bit_exact(engine, pkg, oracle, case):
    """bounds, SFC keys, sort permutation, tree keys: bit-exact against the oracle (itself pinned to the
    reference's kernels by tests/golden)."""
    pos = _cases(pkg)[case]
    st = engine.bh_stages(_dev(pos))
    b = oracle.bh_bounds(pos)
    assert np.array_equal(st["bounds"], b)
    k = oracle.bh_keys(pos, b)
    assert np.array_equal(st["keys"], k)
    _, si = oracle.sort_pairs(k)
    assert np.array_equal(st["sorted_idx"], si)
    assert np.array_equal(st["tree_keys"], oracle.bh_tree_keys(pos, b))


SCHEDULES = {"default": {}, "one_warp_per_group": {"NB_BH_WALK_SPLIT": "1"}}


@pytest.mark.parametrize("schedule", ["default", "one_warp_per_group"])
@pytest.mark.parametrize("case", CASES)
def test_tree_topology_and_accelerations(engine, pkg, oracle, case, schedule, monkeypatch):
    """Both walk schedules -- 2 or 4 warps per group (small systems, the default at these sizes) and one warp per group
    (what systems of 75,000+ bodies run) -- must produce the reference's accepted set: the lane-interaction count equals
    the oracle walk's."""
    import torch
    for k, v in SCHEDULES[schedule].items():
        monkeypatch.setenv(k, v)
    pos = _cases(pkg)[case]
    n = pos.shape[0]
    theta = 0.5
    acc = torch.full((n, 4), 7.0, device="cuda")
    engine.bh_accel(_dev(pos), acc, theta, EPS, pkg.NB_BH_COUNT_WALK)
    torch.cuda.synchronize()
    got = acc.cpu().numpy()
    # topology: identical set of cells (prefix, level, count); monopoles in FP64 on both sides
    c = engine.bh_export_tree()
    b = oracle.bh_bounds(pos)
    _, si = oracle.sort_pairs(oracle.bh_keys(pos, b))
    t = oracle.Tree(pos, b, si)
    oc = t.canonical()
    assert np.array_equal(c["prefix"], oc["prefix"])
    assert np.array_equal(c["level"], oc["level"])
    assert np.array_equal(c["count"], oc["count"])
    assert np.allclose(c["mass"], oc["mass"], rtol=1e-12)
    scale = np.abs(pos[:, :2]).max()
    assert np.abs(c["cx"] - oc["cx"]).max() <= 1e-9 * scale and np.abs(c["cy"] - oc["cy"]).max() <= 1e-9 * scale
    # accelerations: same interaction lists as the oracle walk, FP32 rounding only
    oa, cnt = t.walk(si, theta, EPS)
    assert np.all(got[:, 2:] == 0)
    e = _rel(got, oa.astype(np.float64))
    assert np.median(e) < 2e-6 and np.quantile(e, 0.99) < 2e-5, (np.median(e), np.quantile(e, 0.99), e.max())
    k = engine.bh_last_counters()
    assert abs(k["interactions"] - cnt["interactions"]) <= 1e-4 * cnt["interactions"] + 64
    assert k["status"] == 0


@pytest.mark.parametrize("name", ["bh_disk_n8192", "bh_disk_n16384_t04", "bh_clustered_n8192"])
def test_against_reference_golden(engine, pkg, name):
    """Outputs of the reference's own kernels on B200 (tests/golden): integer stages bit-exact, topology
    identical, accelerations equal up to the reference's float-atomic centre-of-mass noise."""
    import torch
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    pos = g["pos"]
    n = pos.shape[0]
    st = engine.bh_stages(_dev(pos))
    assert np.array_equal(st["bounds"], g["bounds"])
    assert np.array_equal(st["sorted_idx"], g["sorted_idx"])
    assert np.array_equal(st["keys"][g["sorted_idx"]], g["sorted_keys"])
    acc = torch.zeros(n, 4, device="cuda")
    engine.bh_accel(_dev(pos), acc, float(g["theta"]), float(g["eps"]))
    torch.cuda.synchronize()
    c = engine.bh_export_tree()
    assert np.array_equal(c["prefix"], g["cell_prefix"])
    assert np.array_equal(c["level"], g["cell_level"])
    assert np.array_equal(c["count"], g["cell_count"])
    e = _rel(acc.cpu().numpy(), g["acc"].astype(np.float64))
    assert np.median(e) < 5e-6 and np.quantile(e, 0.99) < 1e-3, (np.median(e), np.quantile(e, 0.99), e.max())


def test_theta_error_distribution_matches_oracle(engine, pkg, oracle):
    """theta-error against the FP64 direct sum: same distribution as the reference algorithm (oracle)."""
    import torch
    n = 16384
    pos = pkg.ics.uniform_disk(n, seed=41)
    d = oracle.direct_f64(pos, 2, EPS)
    for theta in (0.3, 0.5, 0.8):
        acc = torch.zeros(n, 4, device="cuda")
        engine.bh_accel(_dev(pos), acc, theta, EPS)
        torch.cuda.synchronize()
        e_gpu = _rel(acc.cpu().numpy(), d)
        e_ora = _rel(oracle.bh_accel(pos, theta, EPS)["acc"], d)
        for q in (0.5, 0.99):
            assert abs(np.quantile(e_gpu, q) - np.quantile(e_ora, q)) <= 0.02 * np.quantile(e_ora, q) + 1e-7
        assert e_gpu.max() <= 1.05 * e_ora.max() + 1e-6


def test_small_theta_converges_to_direct_sum(engine, pkg, oracle):
    import torch
    pos = pkg.ics.uniform_disk(4096, seed=42)
    acc = torch.zeros(4096, 4, device="cuda")
    engine.bh_accel(_dev(pos), acc, 1e-4, EPS)
    torch.cuda.synchronize()
    d = oracle.direct_f64(pos, 2, EPS)
    assert _rel(acc.cpu().numpy(), d).max() < 5e-5


@pytest.mark.parametrize("schedule", ["default", "one_warp_per_group"])
def test_group_slices_tile_the_full_result(engine, pkg, schedule, monkeypatch):
    """Any partition of the groups reproduces the full evaluation bit for bit (given the same number of warps per group)."""
    import torch
    for k, v in SCHEDULES[schedule].items():
        monkeypatch.setenv(k, v)
    n = 10000
    pos = _dev(pkg.ics.uniform_disk(n, seed=43))
    full = torch.zeros(n, 4, device="cuda")
    engine.bh_accel(pos, full, 0.5, EPS)
    parts = torch.full((n, 4), -1.0, device="cuda")
    groups = (n + 31) // 32
    cuts = [0, 1, 100, 101, groups - 1, groups]
    for a, b in zip(cuts[:-1], cuts[1:]):
        engine.bh_accel_groups(pos, parts, 0.5, EPS, a, b - a)
    torch.cuda.synchronize()
    assert torch.equal(full, parts)


def test_deterministic_and_single_body(engine, pkg):
    import torch
    pos = _dev(pkg.ics.clustered(20000, seed=44, sigma=50.0))
    a = torch.zeros(20000, 4, device="cuda")
    b = torch.zeros(20000, 4, device="cuda")
    engine.bh_accel(pos, a, 0.5, EPS)
    engine.bh_accel(pos, b, 0.5, EPS)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    one = _dev(np.array([[1, 2, 0, 3]], np.float32))
    acc = torch.full((1, 4), 5.0, device="cuda")
    engine.bh_accel(one, acc, 0.5, EPS)
    torch.cuda.synchronize()
    assert torch.all(acc == 0)


def test_accumulate_bounds_flag(pkg, oracle):
    """NB_BH_ACCUMULATE_BOUNDS reproduces the reference's never-reset d_bounds (running union)."""
    big = pkg.ics.uniform_disk(4096, seed=45, radius=30000.0)
    small = pkg.ics.uniform_disk(4096, seed=46, radius=1000.0)
    eng = pkg.Engine(0, 4096)
    b1 = eng.bh_stages(_dev(big), pkg.NB_BH_ACCUMULATE_BOUNDS)["bounds"]
    b2 = eng.bh_stages(_dev(small), pkg.NB_BH_ACCUMULATE_BOUNDS)["bounds"]
    assert np.array_equal(b1, oracle.bh_bounds(big))
    assert np.array_equal(b2, oracle.bh_bounds(small, prev=b1)) and np.array_equal(b2, b1)
    b3 = eng.bh_stages(_dev(small))["bounds"]
    assert np.array_equal(b3, oracle.bh_bounds(small))
    eng.close()


def test_host_entry_point(engine, pkg, oracle):
    pos = pkg.ics.uniform_disk(8192, seed=47)
    got = engine.bh_accel_host(pos, 0.5, EPS)
    o = oracle.bh_accel(pos, 0.5, EPS)
    assert np.median(_rel(got, o["acc"].astype(np.float64))) < 2e-6


def test_large_n_properties(pkg, oracle):
    """N = 1,048,576 (BASELINE.json size): sortedness of the key array, permutation validity, stability,
    and accelerations of a sample of bodies against the FP64 direct sum."""
    import torch
    n = 1 << 20
    eng = pkg.Engine(0, n)
    pos = pkg.ics.uniform_disk(n, seed=1)
    dpos = _dev(pos)
    st = eng.bh_stages(dpos)
    si = st["sorted_idx"]
    assert np.array_equal(np.sort(si), np.arange(n, dtype=np.int32))
    sk = st["keys"][si]
    assert np.all(sk[1:] >= sk[:-1])
    ties = sk[1:] == sk[:-1]
    assert np.all(si[1:][ties] > si[:-1][ties])          # stable: ties keep index order
    acc = torch.zeros(n, 4, device="cuda")
    eng.bh_accel(dpos, acc, 0.5, EPS)
    torch.cuda.synchronize()
    sample = np.arange(0, n, n // 512)[:512]
    got = acc.cpu().numpy()[sample]
    ref = np.concatenate([oracle.direct_f64(pos, 2, EPS, t0=int(t), nt=1) for t in sample])
    e = _rel(got, ref)
    assert np.median(e) < 2e-3 and e.max() < 0.05, (np.median(e), e.max())
    eng.close()


def test_sixteen_million_bodies_two_galaxies(pkg, oracle):
    """BASELINE.json config 5 size: N = 16,777,216 two-disc merger, theta = 0.7.  The whole pipeline runs in the
    handle's scratch (< 3 GB), the permutation is valid and a sample of bodies agrees with the FP64 direct sum
    within the theta-error of the method."""
    import torch
    n = 1 << 24
    pos = pkg.ics.two_galaxies(n)
    eng = pkg.Engine(0, n)
    dpos = _dev(pos)
    acc = torch.zeros(n, 4, device="cuda")
    eng.bh_accel(dpos, acc, 0.7, EPS, pkg.NB_BH_COUNT_WALK)
    torch.cuda.synchronize()
    k = eng.bh_last_counters()
    assert k["status"] == 0 and 0 < k["nodes"] < n
    sample = np.arange(0, n, n // 48)[:48]
    got = acc[torch.from_numpy(sample).cuda()].cpu().numpy()
    ref = np.concatenate([oracle.direct_f64(pos, 2, EPS, t0=int(t), nt=1) for t in sample])
    e = _rel(got, ref)
    assert np.median(e) < 5e-3 and e.max() < 0.1, (np.median(e), e.max())
    info = eng.device_info()
    eng.close()


def test_call_sequence_keeps_no_stale_state(pkg):
    """One engine through a random sequence of calls -- sizes on both sides of the graph-replay limit and of the
    engine's current capacity (the scratch grows), repeated arguments (graph replay), changed theta / softening /
    group ranges, the legacy default stream and a private one, stage queries in between -- must give, call for call,
    what a fresh engine gives for the same arguments (the pipeline is deterministic, so bit for bit)."""
    import torch
    rng = np.random.default_rng(2024)
    sizes = [33, 5000, 70000, 300000, 8192, 262144, 262145, 1000, 300000, 70000]
    side = torch.cuda.Stream()
    eng = pkg.Engine(0, 4096)
    pools = {n: _dev(pkg.ics.clustered(n, seed=100 + k, sigma=200.0) if k % 2 else pkg.ics.uniform_disk(n, seed=100 + k))
             for k, n in enumerate(sorted(set(sizes)))}
    for it, n in enumerate(sizes * 2):
        pos = pools[n]
        theta = float(rng.choice([0.3, 0.5, 0.7]))
        eps = float(rng.choice([1.01, 0.25]))
        groups = (n + 31) // 32
        g0 = int(rng.integers(0, groups)) if it % 3 == 1 else 0
        gc = int(rng.integers(0, groups - g0 + 1)) if it % 3 == 1 else groups
        stream = side if it % 4 == 2 else (0 if it % 4 == 3 else None)
        acc = torch.full((n, 4), -3.0, device="cuda")
        want = torch.full((n, 4), -3.0, device="cuda")
        torch.cuda.synchronize()
        for rep in range(3 if it % 2 else 1):     # repeats replay the captured graph where one exists
            eng.bh_accel_groups(pos, acc, theta, eps, g0, gc, stream=stream)
        torch.cuda.synchronize()
        fresh = pkg.Engine(0, n)
        fresh.bh_accel_groups(pos, want, theta, eps, g0, gc)
        torch.cuda.synchronize()
        fresh.close()
        assert torch.equal(acc, want), (it, n, theta, eps, g0, gc)
        if it % 5 == 0:
            st = eng.bh_stages(pos)
            assert np.all(np.diff(st["keys"][st["sorted_idx"]].astype(np.int64)) >= 0)
    eng.close()


def test_dense_system_same_result_on_every_sort_path(pkg):
    """A system with thousands of bodies inside one level-20 cell overflows the five-pass sort's repair block: the
    first evaluation takes the conditional full sort on the device, the following ones are asked for eight plain
    passes by the host.  Every path must give the same tree and the same accelerations, at a graph-replayed size and
    at a plainly launched one."""
    import torch
    for n in (20000, 300000):
        rng = np.random.default_rng(n)
        pos_h = pkg.ics.uniform_disk(n, seed=43)
        pos_h[7000:13000, :2] = pos_h[6999, :2] + rng.uniform(-4e-3, 4e-3, (6000, 2)).astype(np.float32)
        pos = _dev(pos_h)
        eng = pkg.Engine(0, n)
        first, first_tree = None, None
        for rep in range(4):
            acc = torch.zeros(n, 4, device="cuda")
            eng.bh_accel(pos, acc, 0.5, EPS)
            torch.cuda.synchronize()
            tree = eng.bh_export_tree()
            if first is None:
                first, first_tree = acc, tree
            else:
                assert torch.equal(acc, first), rep
                assert all(np.array_equal(tree[k], first_tree[k]) for k in ("prefix", "level", "count"))
        eng.close()


@pytest.mark.parametrize("schedule", ["default", "one_warp_per_group"])
@pytest.mark.parametrize("case", ["disk_8192", "clustered_8192", "duplicates_3000", "dense_knot_6000"])
def test_full_queue_finishes_depth_first(engine, pkg, oracle, case, schedule, monkeypatch):
    """A group whose node queue fills up finishes its walk in depth-first order instead of dropping nodes (the
    reference's 64-entry stack overflows silently, src/BarnesHutKernels.cu:388).  NB_BH_WALK_QCAP lowers the threshold
    so that nearly every group takes that path: same interaction count, same accelerations up to the order of the
    FP32 sum, and the status bit says that it happened."""
    import torch
    for k, v in SCHEDULES[schedule].items():
        monkeypatch.setenv(k, v)
    pos = _cases(pkg)[case]
    n = pos.shape[0]
    dpos = _dev(pos)
    want = torch.zeros(n, 4, device="cuda")
    engine.bh_accel(dpos, want, 0.5, EPS, pkg.NB_BH_COUNT_WALK)
    k0 = engine.bh_last_counters()
    assert k0["status"] == 0
    for qcap in (0, 8, 40):
        monkeypatch.setenv("NB_BH_WALK_QCAP", str(qcap))
        got = torch.full((n, 4), 3.0, device="cuda")
        engine.bh_accel(dpos, got, 0.5, EPS, pkg.NB_BH_COUNT_WALK)
        k1 = engine.bh_last_counters()
        assert k1["status"] == 1 and k1["interactions"] == k0["interactions"], (qcap, k0, k1)
        e = _rel(got.cpu().numpy(), want.cpu().numpy().astype(np.float64))
        assert e.max() < 1e-4 and np.median(e) < 2e-6, (qcap, np.median(e), e.max())
        plain = torch.full((n, 4), 3.0, device="cuda")          # the uninstrumented kernel (graph replay at this size)
        engine.bh_accel(dpos, plain, 0.5, EPS)
        torch.cuda.synchronize()
        assert torch.equal(plain, got)
    monkeypatch.delenv("NB_BH_WALK_QCAP")
    again = torch.zeros(n, 4, device="cuda")
    engine.bh_accel(dpos, again, 0.5, EPS)
    torch.cuda.synchronize()
    assert torch.equal(again, want) and engine.bh_last_counters()["status"] == 0


def test_cold_collapse_one_million(pkg, oracle):
    """BASELINE.json config 4 as specified: a genuinely clustered 1,048,576-body state (ics.cold_collapse_late: knots
    three decades apart in width over a diffuse disc), theta = 0.5.  Checked: permutation validity, sortedness and
    stability of the SFC order, the tree is 10+ levels deeper than the uniform disc's, no group left the batched walk
    (queue depth stays under the threshold), which sort path ran, and 512 sampled bodies against the FP64 direct sum
    within the theta-error of the method."""
    import torch
    n = 1 << 20
    pos = pkg.ics.cold_collapse_late(n, seed=2)
    eng = pkg.Engine(0, n)
    dpos = _dev(pos)
    st = eng.bh_stages(dpos)
    si = st["sorted_idx"]
    assert np.array_equal(np.sort(si), np.arange(n, dtype=np.int32))
    sk = st["keys"][si]
    assert np.all(sk[1:] >= sk[:-1])
    ties = sk[1:] == sk[:-1]
    assert np.all(si[1:][ties] > si[:-1][ties])
    acc = torch.zeros(n, 4, device="cuda")
    eng.bh_accel(dpos, acc, 0.5, EPS, pkg.NB_BH_COUNT_WALK)
    torch.cuda.synchronize()
    k = eng.bh_last_counters()
    info = eng.bh_build_info()
    assert k["status"] == 0 and k["max_stack"] <= 896, k
    assert 0 < info["nodes"] < n and 20 <= info["deepest_level"] <= 32, info
    work = eng.bh_group_work(n).astype(np.int64)             # records each target of a group summed
    assert int(work.sum()) * 32 == k["interactions"]
    blocks = work.reshape(-1, 4).sum(axis=1)                  # four consecutive groups share a thread block
    imbalance = {"group_max_over_mean": float(work.max() / work.mean()), "block_max_over_mean": float(blocks.max() / blocks.mean()),
                 "group_p99_over_mean": float(np.quantile(work, 0.99) / work.mean())}
    sample = np.arange(0, n, n // 512)[:512]
    got = acc.cpu().numpy()[sample]
    ref = np.concatenate([oracle.direct_f64(pos, 2, EPS, t0=int(t), nt=1) for t in sample])
    e = _rel(got, ref)
    assert np.median(e) < 2e-3 and e.max() < 0.05, (np.median(e), e.max())
    again = torch.zeros(n, 4, device="cuda")       # second evaluation: whichever sort path the host now asks for
    eng.bh_accel(dpos, again, 0.5, EPS)
    torch.cuda.synchronize()
    assert torch.equal(again, acc)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        import json
        with open(os.path.join(out, "cold_collapse_1m.json"), "w") as f:
            json.dump({"counters": k, "build": info, "walk_imbalance": imbalance, "theta_error_vs_fp64": {"median": float(np.median(e)),
                       "p99": float(np.quantile(e, 0.99)), "max": float(e.max())}}, f, indent=1)
    eng.close()
