"""Target-slice sharding over the GPUs of one box.

The reference is single-GPU (no NCCL/MPI anywhere); this is new plumbing around the same
force path.  One process per GPU; rank r owns bodies [begin, end) (positions replicated, velocities and
accelerations slice-local).  Per force evaluation ONE all-gather of the owned position slices (in place,
NCCL over NVLink through torch.distributed), then each rank evaluates only its targets:
  direct       -> Engine.direct_accel_slice(pos_full, begin, count, acc_slice)
  Barnes-Hut   -> the tree is rebuilt redundantly from the gathered positions, the walk is sharded by
                  32-target groups of the sorted order: Engine.bh_accel_groups(pos_full, acc_full, g0, g1-g0)
No other collective exists on the path.  The functions below are backend-agnostic (gloo on CPU in the
tests, nccl on GPUs in bench.py).
"""


def slice_bounds(n, world, rank):
    """Bodies owned by `rank`: [begin, end).  Slices differ by at most one body and tile [0, n)."""
    return (n * rank) // world, (n * (rank + 1)) // world


def group_bounds(n, world, rank):
    """32-target groups of the sorted order walked by `rank` (Barnes-Hut): [g0, g1)."""
    groups = (n + 31) // 32
    return (groups * rank) // world, (groups * (rank + 1)) // world


def gather_positions(pos, n, world, rank, dist=None):
    """All-gather of the owned slices of `pos` ([n,4] tensor, replicated layout) in place.

    Equal slices use all_gather_into_tensor with the input aliasing its own part of the output (NCCL's
    in-place all-gather); ragged slices fall back to one broadcast per owner."""
    if world == 1:
        return pos
    if dist is None:
        import torch.distributed as dist
    begin, end = slice_bounds(n, world, rank)
    if n % world == 0:
        dist.all_gather_into_tensor(pos[:n], pos[begin:end])
    else:  # ragged slices (n % world != 0): one broadcast per owner, still in place
        for r in range(world):
            b, e = slice_bounds(n, world, r)
            dist.broadcast(pos[b:e], src=r)
    return pos


class ShardedLeapfrog:
    """Verlet_DKD (src/IntegrationMethods.cu:96-105) on a target slice: half drift, all-gather, force, kick + half
    drift.  `force(pos_full, begin, count, acc_slice)`, `drift(pos_full, vel, begin, count, dt)` and
    `kick_drift(pos_full, vel, acc, begin, count, dt)` are supplied by the caller (Engine methods on GPUs)."""

    def __init__(self, n, world, rank, force, drift, kick_drift, dist=None):
        self.n, self.world, self.rank = n, world, rank
        self.begin, self.end = slice_bounds(n, world, rank)
        self.count = self.end - self.begin
        self.force, self.drift, self.kick_drift, self.dist = force, drift, kick_drift, dist

    def step(self, pos, vel, acc, dt):
        self.drift(pos, vel, self.begin, self.count, dt)
        gather_positions(pos, self.n, self.world, self.rank, self.dist)
        self.force(pos, self.begin, self.count, acc)
        self.kick_drift(pos, vel, acc, self.begin, self.count, dt)
