"""nb_bh_step_groups: the stepper's kick + drift fused into the Barnes-Hut walk, on one GPU (in place) and in the
sharded form (rows in sorted order, scattered back by the next call).  The fused arithmetic is the reference's, so a
trajectory must equal, BIT FOR BIT, the one the separate kernels give (nb_bh_accel + nb_integrate_substep driven by the
mirrored IntegrationMethods classes, themselves checked against the reference's expressions in test_integrators_gpu)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = 1.01
NAMES = ["Euler_Symplectic_KD", "Euler_Symplectic_DK", "Verlet_DKD", "ForestRuth", "PEFRL"]


def _state(pkg, n, seed, vz=False):
    import torch
    pos_h = pkg.ics.uniform_disk(n, seed=seed)
    rng = np.random.default_rng(seed)
    vel_h = np.zeros((n, 4), np.float32)
    vel_h[:, :2] = rng.standard_normal((n, 2)).astype(np.float32) * 2.0e4
    if vz:                       # an R2 run whose bodies carry v_z: the reference's kernels drift z all the same
        pos_h[:, 2] = rng.standard_normal(n).astype(np.float32) * 100.0
        vel_h[:, 2] = rng.standard_normal(n).astype(np.float32) * 50.0
    return torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()


def _classic(pkg, engine, name, pos, vel, dt, steps):
    sysm = pkg.integrators.ShardedSystem(engine, pos.clone(), vel.clone(), "bh", 2, EPS, 0.5)
    integ = getattr(pkg.IntegrationMethods, name)()
    integ.setSystem(sysm)
    for _ in range(steps):
        integ.integrate(dt)
    return sysm.pos, sysm.vel


@pytest.mark.parametrize("name", NAMES)
def test_fused_in_place_equals_separate_kernels(engine, pkg, name):
    import torch
    n, dt, steps = 20000, 1.0e-3, 3
    pos, vel = _state(pkg, n, 7, vz=(name == "Verlet_DKD"))
    want_p, want_v = _classic(pkg, engine, name, pos, vel, dt, steps)
    p, v = pos.clone(), vel.clone()
    fused = pkg.integrators.FusedBarnesHut(pkg.Engine(0, n), p, v, name, 0.5, EPS)
    for _ in range(steps):
        fused.step(dt)
    torch.cuda.synchronize()
    assert torch.equal(p, want_p) and torch.equal(v, want_v)
    assert fused.force_evaluations == steps * len(pkg.integrators.stepper_program(name, dt)[1])


@pytest.mark.parametrize("name", ["Verlet_DKD", "PEFRL"])
def test_fused_step_above_the_eager_capture_size(pkg, name):
    """Above 262,144 bodies an argument set is run directly the first time and captured into a graph only when it
    returns: four steps cross first sighting, capture and replay, and must still equal the separate kernels."""
    import torch
    n, dt, steps = 300000, 1.0e-3, 4
    eng = pkg.Engine(0, n)
    pos, vel = _state(pkg, n, 17)
    want_p, want_v = _classic(pkg, eng, name, pos, vel, dt, steps)
    p, v = pos.clone(), vel.clone()
    fused = pkg.integrators.FusedBarnesHut(pkg.Engine(0, n), p, v, name, 0.5, EPS)
    for _ in range(steps):
        fused.step(dt)
    torch.cuda.synchronize()
    assert torch.equal(p, want_p) and torch.equal(v, want_v)
    # arguments that change on every call (a varying dt) are never captured and stay correct
    q, w = pos.clone(), vel.clone()
    sysm = pkg.integrators.ShardedSystem(eng, pos.clone(), vel.clone(), "bh", 2, EPS, 0.5)
    integ = getattr(pkg.IntegrationMethods, name)()
    integ.setSystem(sysm)
    varying = pkg.integrators.FusedBarnesHut(pkg.Engine(0, n), q, w, name, 0.5, EPS)
    for k in range(3):
        integ.integrate(dt * (1 + k))
        varying.step(dt * (1 + k))
    torch.cuda.synchronize()
    assert torch.equal(q, sysm.pos) and torch.equal(w, sysm.vel)
    eng.close()


@pytest.mark.parametrize("planar", [False, True])
@pytest.mark.parametrize("name,world,n", [("Verlet_DKD", 3, 20000), ("PEFRL", 2, 8192), ("ForestRuth", 4, 5000),
                                          ("Euler_Symplectic_DK", 8, 12345), ("Euler_Symplectic_KD", 2, 33),
                                          ("Verlet_DKD", 2, 270000)])
def test_sharded_rows_equal_separate_kernels(engine, pkg, name, world, n, planar):
    """`world` ranks emulated on one GPU (one engine and one replica of pos / vel each; the all-gather is a
    concatenation of the ranks' rows): after every step each replica equals the single-GPU trajectory bit for bit,
    for even and ragged group counts, with v_z carried through -- or, `planar`, with v_z = 0 and NB_BH_PLANAR, which scatters
    the rows without reading the old state back.  (Sizes at which one GPU and a rank choose the same
    number of warps per group; a different choice changes the order of a target's FP32 sum -- profiles/multigpu_check.py
    pins it with NB_BH_WALK_SPLIT=1 for its two-rank comparison.)"""
    import torch
    I = pkg.integrators
    dt, steps = 1.0e-3, (2 if n <= 262144 else 3)      # above 262,144 the sharded form is launched directly, never replayed
    pos, vel = _state(pkg, n, 11, vz=not planar)
    flags = pkg.NB_BH_PLANAR if planar else 0
    want_p, want_v = _classic(pkg, engine if n <= 262144 else pkg.Engine(0, n), name, pos, vel, dt, steps)
    groups = (n + 31) // 32
    ranks = []
    for r in range(world):
        g0, g1 = pkg.shard.group_bounds(n, world, r)
        ranks.append(dict(eng=pkg.Engine(0, n), pos=pos.clone(), vel=vel.clone(), g0=g0, g1=g1,
                          rows_out=torch.full((32 * max(g1 - g0, 1), 4), float("nan"), device="cuda")))
    rows_all = torch.full((32 * groups, 4), float("nan"), device="cuda")
    pending = None
    for _ in range(steps):
        pre, prog = I.stepper_program(name, dt)
        for k, (kick, drift) in enumerate(prog):
            for rk in ranks:
                rk["eng"].bh_step_groups(rk["pos"], rk["vel"], 0.5, EPS, dt, rk["g0"], rk["g1"] - rk["g0"],
                                         pre_drift=pre if k == 0 else (0, 0.0), kick=kick, drift=drift,
                                         rows_in=rows_all if pending is not None else None,
                                         prev_drift=pending if pending is not None else (0, 0.0), rows_out=rk["rows_out"],
                                         flags=flags)
            torch.cuda.synchronize()
            for rk in ranks:                                  # the all-gather
                rows_all[32 * rk["g0"]:32 * rk["g1"]] = rk["rows_out"][:32 * (rk["g1"] - rk["g0"])]
            pending = drift
    for rk in ranks:                                          # materialize
        rk["eng"].bh_step_groups(rk["pos"], rk["vel"], 0.5, EPS, dt, 0, 0, rows_in=rows_all, prev_drift=pending, flags=flags)
    torch.cuda.synchronize()
    for rk in ranks:
        assert torch.equal(rk["pos"], want_p) and torch.equal(rk["vel"], want_v)
        rk["eng"].close()


def test_fused_stats_match_simulation_stats(engine, pkg):
    """The per-group sums of the epilogue add up to SimulationStats' moment and kinetic energy (nb_stats, itself the
    reference's formulae) of the state after the step."""
    import torch
    n, dt = 30000, 1.0e-3
    pos, vel = _state(pkg, n, 13)
    fused = pkg.integrators.FusedBarnesHut(pkg.Engine(0, n), pos, vel, "Verlet_DKD", 0.5, EPS, stats=True)
    fused.step(dt)
    m, k = fused.step(dt, stats=True)
    want = engine.stats(pos, vel, 2)
    assert abs(m - want[0]) <= 1e-12 * abs(want[0]) and abs(k - want[1]) <= 1e-12 * abs(want[1]), (m, k, want)


def test_fused_step_rejects_what_it_cannot_do(engine, pkg):
    import torch
    I = pkg.integrators
    pos, vel = _state(pkg, 1000, 17)
    with pytest.raises(pkg.NbodyError):
        engine.bh_step_groups(pos, vel, 0.5, EPS, 1e-3, 0, 32, kick=(I.KICK_MEAN_F64, 0.0))
    with pytest.raises(pkg.NbodyError):
        engine.bh_step_groups(pos, vel, 0.5, EPS, 1e-3, 0, 32, drift=(I.DRIFT_VV_F64, 0.0))
    with pytest.raises(pkg.NbodyError):
        engine.bh_step_groups(pos, vel, 0.5, EPS, 1e-3, 0, 33)            # 32 groups only
    fresh = pkg.Engine(0, 1000)
    with pytest.raises(pkg.NbodyError):                                    # rows without a previous evaluation
        fresh.bh_step_groups(pos, vel, 0.5, EPS, 1e-3, 0, 32, rows_in=torch.zeros(1024, 4, device="cuda"))
    fresh.close()
    with pytest.raises(ValueError):
        I.stepper_program("Velocity_Verlet", 1e-3)


# ---- the direct sum: nb_direct_step_slice -------------------------------------------------------------------------------

def _classic_direct(pkg, engine, name, pos, vel, dim, dt, steps, flags):
    sysm = pkg.integrators.ShardedSystem(engine, pos.clone(), vel.clone(), "direct", dim, EPS, direct_flags=flags)
    integ = getattr(pkg.IntegrationMethods, name)()
    integ.setSystem(sysm)
    for _ in range(steps):
        integ.integrate(dt)
    return sysm.pos, sysm.vel


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("n,dim,ordered", [(4096, 2, False), (5000, 3, True), (100000, 3, False)])
def test_fused_direct_equals_separate_kernels(engine, pkg, name, n, dim, ordered):
    """Kick and drift fused into the tile-classification and partial-sum kernels of the direct sum: the trajectory equals the
    separate kernels' bit for bit -- below 65,536 bodies through the replayed CUDA graph, above it through plain launches,
    in the default (split, hoisted) mode and in NB_DIRECT_ORDERED."""
    import torch
    steps, dt = 3, 1.0e-3
    pos, vel = _state(pkg, n, 23, vz=(dim == 3))
    flags = pkg.NB_DIRECT_ORDERED if ordered else pkg.NB_DIRECT_AUTO
    want_p, want_v = _classic_direct(pkg, engine, name, pos, vel, dim, dt, steps, flags)
    p, v = pos.clone(), vel.clone()
    fused = pkg.integrators.FusedDirect(pkg.Engine(0, n), p, v, name, dim, EPS, flags=flags)
    for _ in range(steps):
        fused.step(dt)
    torch.cuda.synchronize()
    assert torch.equal(p, want_p) and torch.equal(v, want_v)


def test_fused_direct_slices_and_stats(engine, pkg):
    """Two slices of one system stepped by two FusedDirect objects (the sharded decomposition, with the all-gather a no-op
    because both work on the same array) reproduce the one-slice trajectory, and the per-block sums of the epilogue are
    SimulationStats' moment and kinetic energy."""
    import torch
    n, dt = 6000, 1.0e-3
    pos, vel = _state(pkg, n, 29)
    want_p, want_v = _classic_direct(pkg, engine, "Verlet_DKD", pos, vel, 2, dt, 1, pkg.NB_DIRECT_AUTO)
    whole = pkg.integrators.FusedDirect(pkg.Engine(0, n), pos.clone(), vel.clone(), "Verlet_DKD", 2, EPS, stats=True)
    m, k = whole.step(dt, stats=True)
    torch.cuda.synchronize()
    assert torch.equal(whole.pos, want_p) and torch.equal(whole.vel, want_v)
    want = engine.stats(whole.pos, whole.vel, 2)
    assert abs(m - want[0]) <= 1e-11 * abs(want[0]) and abs(k - want[1]) <= 1e-11 * abs(want[1]), (m, k, want)   # FP64 sums, other order
    # a slice alone: drift first (own kernel), then the fused force + kick + drift on [a, b)
    a, b = 1000, 4321
    p2, v2 = pos.clone(), vel.clone()
    I = pkg.integrators
    engine.integrate_substep(p2, v2, v2, 0, n, dt, 0, 0.0, I.DRIFT_HALF_F64, 0.0)
    vs = v2[a:b].contiguous()
    engine.direct_step_slice(p2, vs, 2, EPS, dt, a, b - a, kick=(I.KICK_F32, 0.0), drift=(I.DRIFT_HALF_F64, 0.0))
    torch.cuda.synchronize()
    assert torch.equal(p2[a:b], want_p[a:b]) and torch.equal(vs, want_v[a:b])


def test_rows_stored_through_peer_pointers_equal_rows_out(engine, pkg):
    """The exchange inside the walk (n_rows_peers > 0): the epilogue stores each row into every array of a pointer list,
    indexed by the global sorted slot.  On one GPU the 'peers' are three buffers of this process obtained from
    nb_peer_alloc (on a multi-GPU box they are other ranks' buffers mapped with nb_peer_open, tests/test_multigpu_gpu.py):
    each must hold exactly the rows the rows_out form produces, for a group range that starts inside the array."""
    import torch
    I = pkg.integrators
    n, dt = 20000, 1.0e-3
    pos, vel = _state(pkg, n, 31)
    groups = (n + 31) // 32
    g0, gc = 100, groups - 100
    want = torch.full((32 * gc, 4), float("nan"), device="cuda")
    engine.bh_step_groups(pos.clone(), vel.clone(), 0.5, EPS, dt, g0, gc, pre_drift=(I.DRIFT_HALF_F64, 0.0),
                          kick=(I.KICK_F32, 0.0), drift=(I.DRIFT_HALF_F64, 0.0), rows_out=want)
    bufs = [engine.peer_alloc((32 * groups, 4)) for _ in range(3)]
    for b in bufs:
        b.tensor.fill_(float("nan"))
    engine.bh_step_groups(pos.clone(), vel.clone(), 0.5, EPS, dt, g0, gc, pre_drift=(I.DRIFT_HALF_F64, 0.0),
                          kick=(I.KICK_F32, 0.0), drift=(I.DRIFT_HALF_F64, 0.0), rows_peers=[b.ptr for b in bufs])
    torch.cuda.synchronize()
    valid = n - 32 * g0
    for b in bufs:
        t = b.tensor
        assert torch.equal(t[32 * g0:32 * g0 + valid], want[:valid])
        assert torch.isnan(t[:32 * g0]).all()          # slots of groups this call did not walk stay untouched
        b.release()
    with pytest.raises(pkg.NbodyError):
        engine.bh_step_groups(pos, vel, 0.5, EPS, dt, 0, 1, rows_peers=[0])


@pytest.mark.parametrize("name,world,n,steps", [("Verlet_DKD", 2, 20000, 3), ("PEFRL", 3, 8192, 2), ("ForestRuth", 8, 12345, 2),
                                                ("Verlet_DKD", 2, 270000, 4)])
def test_arrival_words_order_the_ranks_without_a_barrier(pkg, name, world, n, steps):
    """nb_bh_step::sync_local: `world` ranks emulated on one GPU, every rank storing its rows into every rank's arrays and
    counting its evaluations into every rank's arrival words; the calls are merely issued in turn on one stream -- no host
    synchronisation, no collective, no copy between them -- and the next call's first kernel waits on the words.  Every
    replica must follow the single-GPU trajectory bit for bit; the word arrays must end at the number of evaluations.
    (n = 270,000: above the sizes at which the sharded form is replayed from a graph.)"""
    import torch
    I = pkg.integrators
    dt = 1.0e-3
    pos, vel = _state(pkg, n, 13, vz=False)
    ref_eng = pkg.Engine(0, n)
    want_p, want_v = _classic(pkg, ref_eng, name, pos, vel, dt, steps)
    groups = (n + 31) // 32
    ranks = []
    for r in range(world):
        g0, g1 = pkg.shard.group_bounds(n, world, r)
        ranks.append(dict(eng=pkg.Engine(0, n), pos=pos.clone(), vel=vel.clone(), g0=g0, g1=g1,
                          rows=[torch.full((32 * groups, 4), float("nan"), device="cuda") for _ in range(2)],
                          words=torch.zeros(16, dtype=torch.int32, device="cuda")))
    sync_ptrs = [rk["words"].data_ptr() for rk in ranks]
    flip, pending, evaluations = 0, None, 0
    for _ in range(steps):
        pre, prog = I.stepper_program(name, dt)
        for k, (kick, drift) in enumerate(prog):
            for r, rk in enumerate(ranks):
                rk["eng"].bh_step_groups(rk["pos"], rk["vel"], 0.5, EPS, dt, rk["g0"], rk["g1"] - rk["g0"],
                                         pre_drift=pre if k == 0 else (0, 0.0), kick=kick, drift=drift,
                                         rows_in=rk["rows"][flip ^ 1] if pending is not None else None,
                                         prev_drift=pending if pending is not None else (0, 0.0),
                                         rows_peers=[q["rows"][flip].data_ptr() for q in ranks], sync=(r, sync_ptrs),
                                         flags=pkg.NB_BH_PLANAR)
            flip ^= 1
            pending = drift
            evaluations += 1
    for r, rk in enumerate(ranks):                            # materialize: waits on the words too
        rk["eng"].bh_step_groups(rk["pos"], rk["vel"], 0.5, EPS, dt, 0, 0, rows_in=rk["rows"][flip ^ 1], prev_drift=pending,
                                 rows_peers=[q["rows"][flip].data_ptr() for q in ranks], sync=(r, sync_ptrs),
                                 flags=pkg.NB_BH_PLANAR)
    torch.cuda.synchronize()
    for rk in ranks:
        assert torch.equal(rk["pos"], want_p) and torch.equal(rk["vel"], want_v)
        w = rk["words"].tolist()
        assert w[:world] == [evaluations] * world and w[8] == evaluations
        assert rk["eng"].bh_last_counters()["status"] & 2 == 0      # no wait timed out
        rk["eng"].close()
    ref_eng.close()
    with pytest.raises(pkg.NbodyError):                      # the words need the pointer list they index
        e = pkg.Engine(0, 64)
        p, v = _state(pkg, 64, 3)
        e.bh_step_groups(p, v, 0.5, EPS, dt, 0, 2, sync=(0, [sync_ptrs[0]]))


def test_replayed_graph_and_pending_rows_survive_other_calls(pkg):
    """Two hazards of cached state.  (i) The fused direct step replays a captured graph at small sizes; a larger call in
    between moves the scratch the graph points to, so the graph must be rebuilt, not replayed.  (ii) rows_in refers to the
    permutation of the handle's last fused evaluation; any other Barnes-Hut evaluation on the handle overwrites it and the
    rows must then be refused rather than scattered with the wrong permutation."""
    import torch
    I = pkg.integrators
    eng = pkg.Engine(0, 1024)
    dt = 1.0e-3
    small_p, small_v = _state(pkg, 4096, 41)
    big_p, big_v = _state(pkg, 150000, 43)
    fresh = pkg.Engine(0, 4096)
    want_p, want_v = small_p.clone(), small_v.clone()
    for _ in range(3):
        fresh.direct_step_slice(want_p, want_v, 2, EPS, dt, 0, 4096, pre_drift=(I.DRIFT_HALF_F64, 0.0), kick=(I.KICK_F32, 0.0),
                                drift=(I.DRIFT_HALF_F64, 0.0))
    p, v = small_p.clone(), small_v.clone()
    kw = dict(pre_drift=(I.DRIFT_HALF_F64, 0.0), kick=(I.KICK_F32, 0.0), drift=(I.DRIFT_HALF_F64, 0.0))
    eng.direct_step_slice(p, v, 2, EPS, dt, 0, 4096, **kw)              # captured
    eng.direct_step_slice(p, v, 2, EPS, dt, 0, 4096, **kw)              # replayed
    eng.direct_step_slice(big_p, big_v, 2, EPS, dt, 0, 150000, **kw)    # grows the partial-sum scratch
    eng.direct_step_slice(p, v, 2, EPS, dt, 0, 4096, **kw)              # same arguments as the captured call
    torch.cuda.synchronize()
    assert torch.equal(p, want_p) and torch.equal(v, want_v)
    # (ii)
    pos, vel = _state(pkg, 8192, 47)
    rows = torch.zeros(8192, 4, device="cuda")
    eng.bh_step_groups(pos, vel, 0.5, EPS, dt, 0, 256, kick=(I.KICK_F32, 0.0), drift=(I.DRIFT_F32, 0.0), rows_out=rows)
    acc = torch.zeros(8192, 4, device="cuda")
    eng.bh_accel(pos, acc, 0.5, EPS)                                     # another evaluation: new permutation buffers
    with pytest.raises(pkg.NbodyError):
        eng.bh_step_groups(pos, vel, 0.5, EPS, dt, 0, 0, rows_in=rows, prev_drift=(I.DRIFT_F32, 0.0))
    eng.close()
    fresh.close()
