"""Seeded synthetic initial conditions.

The reference draws its bodies with std::default_random_engine (src/InitialConditions.cu:111-144),
whose stream differs between standard libraries, so benchmarks and tests use a counter-based generator
(splitmix64 of (seed, stream, index)) that gives the same bodies on every host.  Parameters follow
src/main.cu:26-32 and DiskModel (src/InitialConditions.cu:153-196).
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def uniform01(seed, stream, n):
    """n doubles in [0,1) with 53 random bits, a pure function of (seed, stream, index)."""
    with np.errstate(over="ignore"):
        idx = np.arange(n, dtype=np.uint64)
        base = _splitmix64(np.uint64(seed) * np.uint64(0x100000001B3) + np.uint64(stream))
        bits = _splitmix64(idx ^ base)
    return (bits >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def uniform_disk(n, seed=1, radius=30000.0, mass=70000.0, central_mass=10000.0, z_half_height=0.0,
                 centre=(0.0, 0.0)):
    """float32 [n,4] bodies: phi ~ U[0,2pi), lambda ~ U[0.02,1), pos = radius*lambda*(cos phi, sin phi);
    body 0 is the central mass (src/InitialConditions.cu:157).  z_half_height > 0 gives the R3 variant
    (z ~ U[-h, h]) for the 3D kernel."""
    phi = uniform01(seed, 0, n) * (2.0 * np.pi)
    lam = 0.02 + 0.98 * uniform01(seed, 1, n)
    pos = np.zeros((n, 4), np.float32)
    pos[:, 0] = (centre[0] + radius * lam * np.cos(phi)).astype(np.float32)
    pos[:, 1] = (centre[1] + radius * lam * np.sin(phi)).astype(np.float32)
    if z_half_height > 0:
        pos[:, 2] = ((2.0 * uniform01(seed, 2, n) - 1.0) * z_half_height).astype(np.float32)
    pos[:, 3] = np.float32(mass)
    if n > 0:
        pos[0] = (centre[0], centre[1], 0.0, central_mass)
    return pos


def uniform_box(n, seed=1, centre=(0.0, 0.0), size=(60000.0, 60000.0), space_r3=False, mass_range=(70000.0, 70000.0),
                speed_range=(0.0, 0.0), bulk_velocity=(0.0, 0.0)):
    """UniformBox (src/InitialConditions.cu:78-107): (pos float32 [n,4], vel float32 [n,4]).  In R3 the z coordinate is
    drawn from the y range, as the reference does (:85)."""
    pos = np.zeros((n, 4), np.float32)
    vel = np.zeros((n, 4), np.float32)
    pos[:, 0] = (centre[0] + size[0] * (uniform01(seed, 0, n) - 0.5)).astype(np.float32)
    pos[:, 1] = (centre[1] + size[1] * (uniform01(seed, 1, n) - 0.5)).astype(np.float32)
    if space_r3:
        pos[:, 2] = (centre[1] + size[1] * (uniform01(seed, 2, n) - 0.5)).astype(np.float32)
    pos[:, 3] = (mass_range[0] + (mass_range[1] - mass_range[0]) * uniform01(seed, 3, n)).astype(np.float32)
    phi = uniform01(seed, 4, n) * 6.283185307179586
    v = speed_range[0] + (speed_range[1] - speed_range[0]) * uniform01(seed, 5, n)
    vel[:, 0] = (v * np.cos(phi)).astype(np.float32) + np.float32(bulk_velocity[0])
    vel[:, 1] = (v * np.sin(phi)).astype(np.float32) + np.float32(bulk_velocity[1])
    return pos, vel


def disk_velocities(pos, acc, centre=(0.0, 0.0), space_r3=False):
    """Tangential speeds v = sqrt(|a| r) from one force evaluation, src/InitialConditions.cu:169-192."""
    x = pos[:, 0] - np.float32(centre[0])
    y = pos[:, 1] - np.float32(centre[1])
    z = pos[:, 2] if space_r3 else np.zeros_like(x)
    r = x * x + y * y + np.float32(0.0001)
    a = acc[:, 0] ** 2 + acc[:, 1] ** 2
    if space_r3:
        r = r + z * z
        a = a + acc[:, 2] ** 2
    r = np.sqrt(r, dtype=np.float32)
    a = np.sqrt(a, dtype=np.float32)
    v = np.sqrt(a * r, dtype=np.float32)
    vel = np.zeros((pos.shape[0], 4), np.float32)
    vel[:, 0] = -(v * y / r)
    vel[:, 1] = v * x / r
    vel[:, 2] = v * z / r
    return vel


def two_galaxies(n, seeds=(3, 4), separation=(60000.0, 15000.0), **kw):
    """Two discs of n/2 bodies at (-sx,-sy) and (+sx,+sy) (config 5 of BASELINE.json)."""
    h = n // 2
    a = uniform_disk(h, seed=seeds[0], centre=(-separation[0], -separation[1]), **kw)
    b = uniform_disk(n - h, seed=seeds[1], centre=(separation[0], separation[1]), **kw)
    return np.concatenate([a, b], axis=0)


def clustered(n, seed=5, clusters=8, radius=30000.0, sigma=40.0, mass=70000.0):
    """Deep, imbalanced tree: tight Gaussian clumps (stands in for a late cold-collapse state)."""
    cx = (uniform01(seed, 10, clusters) * 2 - 1) * radius
    cy = (uniform01(seed, 11, clusters) * 2 - 1) * radius
    which = (uniform01(seed, 12, n) * clusters).astype(np.int64) % clusters
    u1 = np.maximum(uniform01(seed, 13, n), 1e-12)
    u2 = uniform01(seed, 14, n)
    rr = sigma * np.sqrt(-2.0 * np.log(u1))
    pos = np.zeros((n, 4), np.float32)
    pos[:, 0] = (cx[which] + rr * np.cos(2 * np.pi * u2)).astype(np.float32)
    pos[:, 1] = (cy[which] + rr * np.sin(2 * np.pi * u2)).astype(np.float32)
    pos[:, 3] = np.float32(mass)
    return pos


def cold_collapse_late(n, seed=2, radius=30000.0, mass=70000.0, clumps=12):
    """A late cold-collapse state (config 4 of BASELINE.json: deep clustered tree, load imbalance): 70 % of the bodies
    sit in `clumps` Gaussian knots whose widths span three decades (5 ... 2000 length units, so the tree is 10 levels
    deeper inside the tightest knot than in the field), populated unequally (weights 1 ... clumps), over a diffuse disc
    holding the remaining 30 %.  Counter-based, same bodies on every host."""
    cx = (uniform01(seed, 20, clumps) * 2 - 1) * 0.6 * radius
    cy = (uniform01(seed, 21, clumps) * 2 - 1) * 0.6 * radius
    sigma = 5.0 * (400.0 ** (np.arange(clumps) / max(clumps - 1, 1)))
    w = np.arange(1, clumps + 1, dtype=np.float64)
    cdf = np.cumsum(w) / w.sum()
    u = uniform01(seed, 22, n)
    in_clump = u < 0.7
    which = np.minimum(np.searchsorted(cdf, u / 0.7), clumps - 1)
    u1 = np.maximum(uniform01(seed, 23, n), 1e-12)
    u2 = uniform01(seed, 24, n)
    rr = sigma[which] * np.sqrt(-2.0 * np.log(u1))
    disc = uniform_disk(n, seed=seed, radius=radius, mass=mass)
    pos = disc.copy()
    pos[:, 0] = np.where(in_clump, (cx[which] + rr * np.cos(2 * np.pi * u2)), disc[:, 0]).astype(np.float32)
    pos[:, 1] = np.where(in_clump, (cy[which] + rr * np.sin(2 * np.pi * u2)), disc[:, 1]).astype(np.float32)
    pos[:, 3] = np.float32(mass)
    return pos


def standard(n, seed=1, centre=(0.0, 0.0)):
    """Standard (src/InitialConditions.cu:13-47, the reference's debug generator): theta ~ U[0, 2 pi),
    r = (Exp(12) + 0.02) * 50000, z = sin(r / 2000) * 2000, mass 100, body 0 = the central mass 10^6.
    Returns (pos float32 [n,4], theta float32 [n]); theta feeds standard_velocities."""
    theta = (uniform01(seed, 0, n) * 6.283185307179586).astype(np.float32)
    r = ((-np.log(1.0 - uniform01(seed, 1, n)) / 12.0 + 0.02) * 50000.0).astype(np.float32)
    pos = np.zeros((n, 4), np.float32)
    pos[:, 0] = (centre[0] + r.astype(np.float64) * np.cos(theta.astype(np.float64))).astype(np.float32)
    pos[:, 1] = (centre[1] + r.astype(np.float64) * np.sin(theta.astype(np.float64))).astype(np.float32)
    pos[:, 2] = np.sin(r / np.float32(2000.0), dtype=np.float32) * np.float32(2000.0)
    pos[:, 3] = np.float32(100.0)
    if n > 0:
        pos[0] = (centre[0], centre[1], 0.0, 1000000.0)
    return pos, theta


def standard_velocities(pos, acc, theta, rotation=1.0):
    """The velocity loop of Standard::initialize (src/InitialConditions.cu:49-73): v = sqrt(|F| dist) along
    rotation * (sin theta, -cos theta); the central body stays at rest."""
    d = pos[0, :3] - pos[:, :3]
    dist = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2] + np.float32(0.0001), dtype=np.float32)
    F = np.sqrt(acc[:, 0] * acc[:, 0] + acc[:, 1] * acc[:, 1] + acc[:, 2] * acc[:, 2], dtype=np.float32)
    v = np.sqrt(dist * F, dtype=np.float32)
    vel = np.zeros((pos.shape[0], 4), np.float32)
    rot = np.float32(rotation)
    vel[:, 0] = (rot * v) * np.sin(theta.astype(np.float64)).astype(np.float32)
    vel[:, 1] = (-rot * v) * np.cos(theta.astype(np.float64)).astype(np.float32)
    vel[0] = 0
    return vel
