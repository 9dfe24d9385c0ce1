"""The reference's steppers (src/IntegrationMethods.cu) rebuilt on nb_integrate_substep: every kick/drift kind against
a NumPy restatement of the reference expression (types included), and the order of accuracy of each scheme on a
two-body orbit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = 1.01


def _ulp_close(a, b, ulps=1):
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    tol = ulps * np.spacing(np.maximum(np.abs(a), np.abs(b)).astype(np.float32))
    return np.all(np.abs(a.astype(np.float64) - b.astype(np.float64)) <= tol)


def test_every_kick_and_drift_kind_matches_the_reference_expression(engine, pkg):
    import torch
    I = pkg.integrators
    rng = np.random.default_rng(3)
    n, begin = 1000, 100
    pos = (rng.standard_normal((n + 2 * begin, 4)) * 3e4).astype(np.float32)
    vel = (rng.standard_normal((n, 4)) * 300).astype(np.float32)
    acc = (rng.standard_normal((n, 4)) * 7).astype(np.float32)
    accp = (rng.standard_normal((n, 4)) * 7).astype(np.float32)
    dt = np.float32(0.001)
    d64 = np.float64(dt)
    ck, cd = 1.35120719195966, -0.7024143839193199
    f64 = lambda x: x.astype(np.float64)
    kicks = {
        I.KICK_NONE: lambda v, a, ap: v,
        I.KICK_F32: lambda v, a, ap: (f64(v) + d64 * f64(a)).astype(np.float32),
        I.KICK_HALF_F64: lambda v, a, ap: (f64(v) + (d64 * 0.5) * f64(a)).astype(np.float32),
        I.KICK_COEF_F64: lambda v, a, ap: (f64(v) + ck * f64(a)).astype(np.float32),
        I.KICK_DTA_F64: lambda v, a, ap: (f64(v) + f64((dt * a).astype(np.float32)) * ck).astype(np.float32),
        I.KICK_MEAN_F64: lambda v, a, ap: (f64(v) + (0.5 * d64) * f64((a + ap).astype(np.float32))).astype(np.float32),
    }
    drifts = {
        I.DRIFT_NONE: lambda p, v, a: p,
        I.DRIFT_F32: lambda p, v, a: (f64(p) + f64(v) * d64).astype(np.float32),
        I.DRIFT_HALF_F64: lambda p, v, a: (f64(p) + f64(v) * (d64 * 0.5)).astype(np.float32),
        I.DRIFT_COEF_F64: lambda p, v, a: (f64(p) + cd * f64(v)).astype(np.float32),
        I.DRIFT_VDT_F64: lambda p, v, a: (f64(p) + (f64((v * dt).astype(np.float32)) * cd) * 0.5).astype(np.float32),
        I.DRIFT_VV_F64: lambda p, v, a: (f64(p) + (f64((v * dt).astype(np.float32)) + f64(a) * 0.5 * d64 * d64)).astype(np.float32),
    }
    for kk, kf in kicks.items():
        for dk, df in drifts.items():
            dp, dv = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
            engine.integrate_substep(dp, dv, torch.from_numpy(acc).cuda(), begin, n, float(dt), kk, ck, dk, cd,
                                     acc_prev=torch.from_numpy(accp).cuda())
            torch.cuda.synchronize()
            v_ref = vel.copy()
            v_ref[:, :3] = kf(vel[:, :3], acc[:, :3], accp[:, :3])
            p_ref = pos.copy()
            p_ref[begin:begin + n, :3] = df(pos[begin:begin + n, :3], v_ref[:, :3], acc[:, :3])
            assert _ulp_close(dv.cpu().numpy(), v_ref), (kk, dk)
            assert _ulp_close(dp.cpu().numpy(), p_ref), (kk, dk)
            # w components (mass, padding) and bodies outside the slice are untouched
            assert np.array_equal(dp.cpu().numpy()[:, 3], pos[:, 3]) and np.array_equal(dv.cpu().numpy()[:, 3], vel[:, 3])
            assert np.array_equal(dp.cpu().numpy()[:begin], pos[:begin]) and np.array_equal(dp.cpu().numpy()[begin + n:], pos[begin + n:])


def _two_body(pkg, engine, integrator, steps_per_orbit, orbits=1):
    """Equal masses on a circular orbit (softening 0.01 on r^2 = 4): relative energy error after `orbits`."""
    import torch
    m, r, eps = 1000.0, 1.0, 0.01
    d2 = 4 * r * r + eps
    a = m * 2 * r / d2 ** 1.5                       # acceleration magnitude of the reference's 3D law
    v = np.sqrt(a * r)
    pos = np.array([[r, 0, 0, m], [-r, 0, 0, m]], np.float32)
    vel = np.array([[0, v, 0, 0], [0, -v, 0, 0]], np.float32)
    period = 2 * np.pi * r / v
    dt = period / steps_per_orbit
    sysm = pkg.integrators.ShardedSystem(engine, torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda(), "direct", 3, eps)
    integ = integrator()
    integ.setSystem(sysm)

    def energy():
        p, w = sysm.pos.cpu().numpy().astype(np.float64), sysm.vel.cpu().numpy().astype(np.float64)
        ke = 0.5 * m * (w[:, :3] ** 2).sum()
        dd = ((p[0, :3] - p[1, :3]) ** 2).sum() + eps
        return ke - m * m / np.sqrt(dd)
    e0 = energy()
    for _ in range(steps_per_orbit * orbits):
        integ.integrate(dt)
    torch.cuda.synchronize()
    return abs((energy() - e0) / e0), np.linalg.norm(sysm.pos.cpu().numpy()[0, :3] - pos[0, :3]) / r


@pytest.mark.parametrize("name,order", [("Euler_Symplectic_KD", 1), ("Euler_Symplectic_DK", 1), ("Verlet_DKD", 2),
                                        ("Verlet_KDK", 2), ("Velocity_Verlet", 2), ("ForestRuth", 4), ("PEFRL", 4)])
def test_order_of_accuracy_on_a_two_body_orbit(engine, pkg, name, order):
    """Halving dt shrinks the closure error of one orbit by about 2^order (position returns to its start)."""
    cls = getattr(pkg.IntegrationMethods, name)
    _, x1 = _two_body(pkg, engine, cls, 40)
    _, x2 = _two_body(pkg, engine, cls, 80)
    ratio = x1 / max(x2, 1e-12)
    assert x2 < x1
    if order == 1:
        assert ratio > 1.5, (x1, x2)          # symplectic Euler closes a periodic orbit better than its local order
    elif order == 2:
        assert 3.0 < ratio < 5.5, (x1, x2)
    else:
        assert ratio > 9.0 and x2 < 1e-3, (x1, x2)


def test_sharded_system_barnes_hut_leapfrog(engine, pkg):
    """ShardedSystem + Verlet_DKD on the Barnes-Hut method (world 1): one step equals the hand-written sequence."""
    import torch
    n = 8192
    pos_h = pkg.ics.uniform_disk(n, seed=61)
    pos = torch.from_numpy(pos_h).cuda()
    vel = torch.zeros(n, 4, device="cuda")
    sysm = pkg.integrators.ShardedSystem(engine, pos.clone(), vel.clone(), "bh", 2, EPS, 0.5)
    lf = pkg.IntegrationMethods.Verlet_DKD()
    lf.setSystem(sysm)
    lf.integrate(0.001)
    p2, v2, a2 = pos.clone(), vel.clone(), torch.zeros(n, 4, device="cuda")
    engine.verlet_dkd_drift(p2, v2, 0, n, 0.001)
    engine.bh_accel(p2, a2, 0.5, EPS)
    engine.verlet_dkd_kick_drift(p2, v2, a2, 0, n, 0.001)
    torch.cuda.synchronize()
    assert torch.equal(sysm.pos, p2) and torch.equal(sysm.vel, v2)


def test_simulation_stats_match_the_reference_formulae(engine, pkg):
    """nb_stats against a NumPy FP64 restatement of src/SimulationStats.cu:29-90 (pairs counted twice, i = j included,
    epsilon 1e-6), in 2D and 3D, whole system and as the sum of ragged slices."""
    import torch
    n = 3000
    pos = pkg.ics.uniform_disk(n, seed=71, z_half_height=3000.0)
    rng = np.random.default_rng(4)
    vel = (rng.standard_normal((n, 4)) * 200).astype(np.float32)
    p64, v64, m = pos[:, :3].astype(np.float64), vel.astype(np.float64), pos[:, 3].astype(np.float64)
    d2 = ((p64[:, None, :] - p64[None, :, :]) ** 2).sum(-1) + 1e-6
    pe = -(m[:, None] * m[None, :] / np.sqrt(d2)).sum()
    dpos, dvel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    for dim in (2, 3):
        if dim == 3:
            v = np.sqrt((v64[:, :3] ** 2).sum(1))
        else:
            v = np.sqrt((vel[:, 0] * vel[:, 0] + vel[:, 1] * vel[:, 1]).astype(np.float64))
        want = ((v * m).sum(), (0.5 * v * v * m).sum(), pe)
        got = engine.stats(dpos, dvel, dim)
        # 2D: the reference squares and adds in FP32 before the sqrt (:34); the FMA contraction of the device differs
        # from NumPy's two roundings by an FP32 ulp per body
        rt = 1e-12 if dim == 3 else 1e-7
        assert np.allclose(got[:2], want[:2], rtol=rt) and np.isclose(got[2], want[2], rtol=1e-12), (got, want)
        parts = np.zeros(3)
        for a, b in ((0, 1), (1, 1300), (1300, n)):
            parts += np.array(engine.stats(dpos, dvel[a:b].contiguous(), dim, a, b - a))
        assert np.allclose(parts[:2], want[:2], rtol=rt) and np.isclose(parts[2], want[2], rtol=1e-12)


def test_device_initial_conditions_match_the_host_generators(engine, pkg):
    """nb_ic_uniform_ellipsoid / nb_ic_disk_velocities against ics.uniform_disk / ics.disk_velocities (same
    counter-based generator): positions equal up to one FP32 ulp where the device's and the host's double cos/sin
    differ in the last bit, velocities (pure FP32 arithmetic, IEEE on both sides) bit for bit."""
    import torch
    n = 200000
    for zh in (0.0, 3000.0):
        host = pkg.ics.uniform_disk(n, seed=5, z_half_height=zh)
        pos = torch.zeros(n + 10, 4, device="cuda")
        engine.ic_uniform_ellipsoid(pos, 10, n, seed=5, z_half_height=zh)
        torch.cuda.synchronize()
        dev = pos.cpu().numpy()
        assert np.all(dev[:10] == 0)
        d = dev[10:]
        assert np.array_equal(d[:, 2:], host[:, 2:])
        ulp = np.spacing(np.abs(host[:, :2]).astype(np.float32))
        assert np.all(np.abs(d[:, :2] - host[:, :2]) <= ulp) and np.mean(d[:, :2] == host[:, :2]) > 0.9999
    # velocities from a force evaluation, as DiskModel does
    host = pkg.ics.uniform_disk(n, seed=5)
    dpos = torch.from_numpy(host).cuda()
    acc = torch.zeros(n, 4, device="cuda")
    engine.bh_accel(dpos, acc, 0.5, EPS)
    vel = torch.zeros(n, 4, device="cuda")
    engine.ic_disk_velocities(dpos, acc, vel, 0, n)
    torch.cuda.synchronize()
    want = pkg.ics.disk_velocities(host, acc.cpu().numpy())
    assert np.array_equal(vel.cpu().numpy(), want)


def test_device_uniform_box_matches_the_host_generator(engine, pkg):
    """nb_ic_uniform_box against ics.uniform_box: positions and masses bit for bit (explicitly rounded double
    arithmetic on both sides), velocities within one FP32 ulp (double cos/sin of the device vs the host)."""
    import torch
    n = 100000
    for r3 in (False, True):
        kw = dict(seed=11, centre=(500.0, -2500.0), size=(40000.0, 10000.0), mass_range=(10.0, 90000.0),
                  speed_range=(5.0, 250.0), bulk_velocity=(30.0, -4.0))
        hp, hv = pkg.ics.uniform_box(n, space_r3=r3, **kw)
        pos = torch.zeros(n + 7, 4, device="cuda")
        vel = torch.zeros(n, 4, device="cuda")
        engine.ic_uniform_box(pos, 7, n, dim=3 if r3 else 2, vel=vel, **kw)
        torch.cuda.synchronize()
        dp, dv = pos.cpu().numpy(), vel.cpu().numpy()
        assert np.all(dp[:7] == 0) and np.array_equal(dp[7:], hp)
        assert np.all(np.abs(hp[:, 0] - 500.0) <= 20000.0) and np.all(np.abs(hp[:, 1] + 2500.0) <= 5000.0)
        assert (hp[:, 2].std() > 1000.0) == r3
        ulp = np.spacing(np.maximum(np.abs(hv[:, :2]), 1.0).astype(np.float32))
        assert np.all(np.abs(dv[:, :2] - hv[:, :2]) <= ulp) and np.mean(dv[:, :2] == hv[:, :2]) > 0.999
        assert np.all(dv[:, 2:] == 0)
    # empty range and bad arguments
    engine.ic_uniform_box(pos, 0, 0)
    with pytest.raises(RuntimeError):
        engine.ic_uniform_box(pos, 0, 10, dim=4)


def test_device_standard_generator_matches_the_host_twin(engine, pkg):
    """nb_ic_standard / nb_ic_standard_velocities (InitialConditions::Standard, src/InitialConditions.cu:13-76) against
    ics.standard / ics.standard_velocities: masses and the central body bit for bit, positions and velocities within one
    FP32 ulp (double cos / sin / log and FP32 sin of the device against the host's)."""
    import torch
    n = 100000
    host, theta = pkg.ics.standard(n, seed=9, centre=(100.0, -50.0))
    pos = torch.zeros(n + 3, 4, device="cuda")
    engine.ic_standard(pos, 3, n, seed=9, centre=(100.0, -50.0))
    torch.cuda.synchronize()
    dev = pos.cpu().numpy()
    assert np.all(dev[:3] == 0) and np.array_equal(dev[3:, 3], host[:, 3]) and np.array_equal(dev[3], host[0])
    ulp = np.spacing(np.maximum(np.abs(host[:, :3]), 1.0).astype(np.float32))
    assert np.all(np.abs(dev[3:, :3] - host[:, :3]) <= 2 * ulp) and np.mean(dev[3:, :2] == host[:, :2]) > 0.99
    # velocities from a force evaluation on the HOST positions (so that both sides see the same accelerations)
    dpos = torch.from_numpy(host).cuda()
    acc = torch.zeros(n, 4, device="cuda")
    engine.direct_accel(dpos, acc, 3, EPS)
    vel = torch.zeros(n, 4, device="cuda")
    engine.ic_standard_velocities(dpos, acc, vel, 0, n, seed=9, rotation=-1.0)
    torch.cuda.synchronize()
    want = pkg.ics.standard_velocities(host, acc.cpu().numpy(), theta, rotation=-1.0)
    got = vel.cpu().numpy()
    ulp = np.spacing(np.maximum(np.abs(want), 1e-3).astype(np.float32))
    assert np.all(got[0] == 0) and np.all(np.abs(got - want) <= 2 * ulp), np.abs(got - want).max()
