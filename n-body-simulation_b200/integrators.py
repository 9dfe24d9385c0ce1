"""The reference's steppers (namespace IntegrationMethods, src/IntegrationMethods.{cuh,cu}) on the engine.

Same class names and the same sequence of kernels and force evaluations per step; the elementwise kernels are
nb_integrate_substep calls on the rank's target slice, the force evaluation is whatever
`system.force()` does (direct or Barnes-Hut, preceded by the all-gather of positions when sharded).

`system` must provide: engine, pos (replicated [N,4]), vel / acc (slice-local), begin, count, force(), gather().
"""
import os
import numpy as np

KICK_NONE, KICK_F32, KICK_HALF_F64, KICK_COEF_F64, KICK_DTA_F64, KICK_MEAN_F64 = range(6)
DRIFT_NONE, DRIFT_F32, DRIFT_HALF_F64, DRIFT_COEF_F64, DRIFT_VDT_F64, DRIFT_VV_F64 = range(6)


class IntegrationMethods:
    class IntegrationMethod:
        force_evaluations = 1

        def setSystem(self, system):                      # src/IntegrationMethods.cu:7-9
            self.system = system

        def _sub(self, dt, kick=0, kc=0.0, drift=0, dc=0.0, acc_prev=None):
            s = self.system
            s.engine.integrate_substep(s.pos, s.vel, s.acc, s.begin, s.count, dt, kick, kc, drift, dc, acc_prev)

        def _force(self):
            s = self.system
            s.gather()
            s.force()

        def integrate(self, dt):
            raise NotImplementedError

    class Euler_Symplectic_KD(IntegrationMethod):          # :26-35
        def integrate(self, dt):
            self._force()
            self._sub(dt, KICK_F32, 0.0, DRIFT_F32)

    class Euler_Symplectic_DK(IntegrationMethod):          # :58-67
        def integrate(self, dt):
            self._sub(dt, drift=DRIFT_F32)
            self._force()
            self._sub(dt, kick=KICK_F32)

    class Verlet_DKD(IntegrationMethod):                   # :96-105  ("Leapfrog")
        def integrate(self, dt):
            self._sub(dt, drift=DRIFT_HALF_F64)
            self._force()
            self._sub(dt, KICK_F32, 0.0, DRIFT_HALF_F64)

    class Verlet_KDK(IntegrationMethod):                   # :133-146 (declared but not linkable in the reference)
        def __init__(self):
            self.firstCalc = True

        def integrate(self, dt):
            if self.firstCalc:
                self.firstCalc = False
                self._force()
            self._sub(dt, KICK_HALF_F64, 0.0, DRIFT_F32)
            self._force()
            self._sub(dt, kick=KICK_HALF_F64)

    class Velocity_Verlet(IntegrationMethod):              # :168-190
        def __init__(self):
            self.firstCalc = True
            self.prev_acc = None

        def integrate(self, dt):
            import torch
            s = self.system
            if self.prev_acc is None:
                self.prev_acc = torch.zeros_like(s.acc)
            if self.firstCalc:
                self.firstCalc = False
                self._force()
            self._sub(dt, drift=DRIFT_VV_F64)
            self.prev_acc, s.acc = s.acc, self.prev_acc    # std::swap(d_prev_acc, system->device.acc), :184
            self._force()
            self._sub(dt, kick=KICK_MEAN_F64, acc_prev=self.prev_acc)

    class ForestRuth(IntegrationMethod):                   # :197-270, three force evaluations
        force_evaluations = 3
        O = 1.35120719195966

        def integrate(self, dt):
            O = self.O
            self._sub(dt, drift=DRIFT_VDT_F64, dc=O)
            self._force()
            self._sub(dt, KICK_DTA_F64, O, DRIFT_VDT_F64, 1 - O)
            self._force()
            self._sub(dt, KICK_DTA_F64, 1 - 2 * O, DRIFT_VDT_F64, 1 - O)
            self._force()
            self._sub(dt, KICK_DTA_F64, O, DRIFT_VDT_F64, O)

    class PEFRL(IntegrationMethod):                        # :274-374, four force evaluations
        force_evaluations = 4
        E, L, X = 0.1786178958448091, -0.2123418310626054, -0.06626458266981849

        def integrate(self, dt):
            d = float(np.float32(dt))                      # the kernels take `float dt` and promote it (:276)
            Et, Xt, Lt = self.E * d, self.X * d, self.L * d
            mL = (1 + 2 * 0.2123418310626054) / 2 * d
            kt = (1 - 2 * (0.1786178958448091 - 0.06626458266981849)) * d
            self._sub(dt, drift=DRIFT_COEF_F64, dc=Et)
            self._force()
            self._sub(dt, KICK_COEF_F64, mL, DRIFT_COEF_F64, Xt)
            self._force()
            self._sub(dt, KICK_COEF_F64, Lt, DRIFT_COEF_F64, kt)
            self._force()
            self._sub(dt, KICK_COEF_F64, Lt, DRIFT_COEF_F64, Xt)
            self._force()
            self._sub(dt, KICK_COEF_F64, mL, DRIFT_COEF_F64, Et)


class ShardedSystem:
    """State of one rank: replicated positions, slice-local velocities / accelerations, a force method."""

    def __init__(self, engine, pos, vel, method="direct", dim=3, softening=1.01, theta=0.5, world=1, rank=0, dist=None,
                 direct_flags=0):
        import torch
        from . import shard
        self.engine, self.pos, self.vel = engine, pos, vel
        self.n = pos.shape[0]
        self.world, self.rank, self.dist = world, rank, dist
        self.begin, self.end = shard.slice_bounds(self.n, world, rank)
        self.count = self.end - self.begin
        self.method, self.dim, self.eps, self.theta = method, dim, softening, theta
        self.direct_flags = direct_flags
        self.acc = torch.zeros(self.count, 4, device=pos.device)
        self._acc_full = torch.zeros(self.n, 4, device=pos.device) if method == "bh" else None
        self._shard = shard

    def gather(self):
        self._shard.gather_positions(self.pos, self.n, self.world, self.rank, self.dist)

    def force(self):
        if self.method == "direct":
            self.engine.direct_accel_slice(self.pos, self.begin, self.count, self.acc, self.dim, self.eps, self.direct_flags)
        else:
            # Barnes-Hut shards by groups of the sorted order, not by body slice: walk the rank's groups, then
            # exchange so that every rank holds the accelerations of its own bodies
            g0, g1 = self._shard.group_bounds(self.n, self.world, self.rank)
            if self.world > 1:
                self._acc_full.zero_()
            self.engine.bh_accel_groups(self.pos, self._acc_full, self.theta, self.eps, g0, g1 - g0)
            if self.world > 1 and self.n % self.world == 0:
                self.dist.reduce_scatter_tensor(self.acc, self._acc_full)      # each rank keeps its own bodies' rows
            else:
                if self.world > 1:
                    self.dist.all_reduce(self._acc_full)
                self.acc.copy_(self._acc_full[self.begin:self.end])


# Every stepper except Velocity Verlet as  pre-force drift + [force; kick; drift]*  (kind, coefficient) pairs -- the
# kernel sequences of src/IntegrationMethods.cu:26-35, 58-67, 96-105, 245-270, 343-374.  `d` is float(np.float32(dt)).
def stepper_program(name, dt):
    d = float(np.float32(dt))
    if name == "Euler_Symplectic_KD":
        return (DRIFT_NONE, 0.0), [((KICK_F32, 0.0), (DRIFT_F32, 0.0))]
    if name == "Euler_Symplectic_DK":
        return (DRIFT_F32, 0.0), [((KICK_F32, 0.0), (DRIFT_NONE, 0.0))]
    if name == "Verlet_DKD":
        return (DRIFT_HALF_F64, 0.0), [((KICK_F32, 0.0), (DRIFT_HALF_F64, 0.0))]
    if name == "ForestRuth":
        O = IntegrationMethods.ForestRuth.O
        return (DRIFT_VDT_F64, O), [((KICK_DTA_F64, O), (DRIFT_VDT_F64, 1 - O)),
                                    ((KICK_DTA_F64, 1 - 2 * O), (DRIFT_VDT_F64, 1 - O)),
                                    ((KICK_DTA_F64, O), (DRIFT_VDT_F64, O))]
    if name == "PEFRL":
        P = IntegrationMethods.PEFRL
        Et, Xt, Lt = P.E * d, P.X * d, P.L * d
        mL = (1 + 2 * 0.2123418310626054) / 2 * d
        kt = (1 - 2 * (0.1786178958448091 - 0.06626458266981849)) * d
        return (DRIFT_COEF_F64, Et), [((KICK_COEF_F64, mL), (DRIFT_COEF_F64, Xt)), ((KICK_COEF_F64, Lt), (DRIFT_COEF_F64, kt)),
                                      ((KICK_COEF_F64, Lt), (DRIFT_COEF_F64, Xt)), ((KICK_COEF_F64, mL), (DRIFT_COEF_F64, Et))]
    raise ValueError(f"{name} has no fused form (Velocity_Verlet and Verlet_KDK need the previous acceleration)")


class FusedBarnesHut:
    """Barnes-Hut time stepping with the kick and drift fused into the walk (nb_bh_step_groups), on one GPU or sharded.

    Sharded (world > 1): positions AND velocities are replicated on every rank; a rank walks its share of the 32-target
    groups of the sorted order, its epilogue produces (x', y', vx', vy') rows in sorted order, ONE all-gather per force
    evaluation (16 bytes per body) replicates them, and the next evaluation's first kernel scatters them back to body
    order.  That replaces the all-gather of positions plus the N x 16-byte exchange of accelerations of ShardedSystem.
    `pos` / `vel` are current after materialize() (called by step() unless lazy=True)."""

    def __init__(self, engine, pos, vel, integrator="Verlet_DKD", theta=0.5, softening=1.01, world=1, rank=0, dist=None,
                 flags=0, stats=False, p2p=False):
        """p2p (sharded runs on one box): no all-gather call at all -- the walk's epilogue stores each row straight into
        every rank's rows array over NVLink (the arrays are allocated by nb_peer_alloc and mapped by the other ranks);
        two arrays alternate, and consecutive force evaluations are separated by one tiny all-reduce used as a barrier."""
        import torch
        from . import shard
        self.engine, self.pos, self.vel = engine, pos, vel
        self.n = pos.shape[0]
        self.integrator, self.theta, self.eps, self.flags = integrator, theta, softening, flags
        self.world, self.rank, self.dist = world, rank, dist
        if world > 1 and not bool((vel[:, 2] != 0).any().item()):
            self.flags |= 0x200      # NB_BH_PLANAR: v_z = 0 everywhere and the 2D path keeps it so -- z never moves
        self.groups = (self.n + 31) // 32
        self.g0, self.g1 = shard.group_bounds(self.n, world, rank)
        self._bounds = [shard.group_bounds(self.n, world, r) for r in range(world)]
        self.even = self.groups % world == 0
        self.rows_all = self.rows_out = None
        if world > 1:
            self.rows_all = torch.empty(32 * self.groups, 4, device=pos.device)
            self.rows_out = self.rows_all[32 * self.g0:32 * self.g1] if self.even else \
                torch.empty(32 * (self.g1 - self.g0), 4, device=pos.device)
        self.stats_buf = torch.zeros(2 * self.groups, dtype=torch.float64, device=pos.device) if stats else None
        self.pending = None          # ((drift kind, coefficient), dt) of the evaluation whose rows are waiting in rows_all
        self.force_evaluations = 0
        self.p2p = bool(p2p) and 1 < world <= 8      # NB_MAX_PEERS pointers fit in one nb_bh_step
        if self.p2p:
            # every rank maps every other rank's two rows arrays; if any rank cannot (no peer access between the GPUs, no
            # IPC between the processes), all ranks fall back to the all-gather together
            # two rows arrays (alternating) and the 16 arrival words of nb_bh_step::sync_local
            shapes = [(32 * self.groups, 4), (32 * self.groups, 4), (16,)]
            ok, opened = 1, []
            try:
                self._mine = [engine.peer_alloc(sh) for sh in shapes]
                self._mine[2].tensor.zero_()
                mine = [b.handle for b in self._mine]
            except Exception:
                ok, mine, self._mine = 0, [None, None, None], []
            handles = [None] * world
            dist.all_gather_object(handles, mine)
            if ok and all(h[0] is not None for h in handles):
                try:
                    self._peers = [[self._mine[k] if r == rank else engine.peer_open(handles[r][k], shapes[k])
                                    for r in range(world)] for k in range(3)]
                    opened = [b for row in self._peers for b in row if b.opened]
                except Exception:
                    ok = 0
            else:
                ok = 0
            flag = torch.tensor([float(ok)], device=pos.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)      # also: everyone has mapped everyone before the first store
            if flag.item() >= 1.0:
                self._rows = [b.tensor for b in self._mine[:2]]
                self._flip = 0
                self._token = torch.zeros(1, device=pos.device)
                # NB_BH_SYNC=0: one tiny all-reduce as the barrier between evaluations instead of the arrival words
                self._sync = (rank, [b.ptr for b in self._peers[2]]) if os.environ.get("NB_BH_SYNC", "1") != "0" else None
            else:
                for b in opened + list(self._mine):
                    b.release()
                self.p2p = False

    def close(self):
        """Unmaps the other ranks' rows arrays and frees this rank's (p2p); call on every rank before the engine goes."""
        if self.p2p:
            import torch
            torch.cuda.synchronize()
            self.dist.barrier()          # no rank may still be storing into an array that is about to go
            for row in self._peers:
                for b in row:
                    b.release()
            self.p2p = False

    def _exchange(self):
        """What follows a rank's walk: the all-gather of the rows, or (p2p) the barrier behind the peer stores -- unless
        the calls order themselves through the arrival words (nb_bh_step::sync_local), then nothing."""
        if self.p2p:
            if self._sync is None:
                self.dist.all_reduce(self._token)
        else:
            self._gather_rows()

    def _gather_rows(self):
        if self.even:                # in place: every rank's rows already sit at their offset of rows_all
            self.dist.all_gather_into_tensor(self.rows_all, self.rows_out)
        else:                        # ragged group counts: one broadcast per owner
            self.rows_all[32 * self.g0:32 * self.g1].copy_(self.rows_out)
            for r, (a, b) in enumerate(self._bounds):
                self.dist.broadcast(self.rows_all[32 * a:32 * b], src=r)

    def _group(self, dt, pre, kick, drift, stats):
        e = self.engine
        sharded = self.world > 1
        rows_in, prev = (self.rows_all, self.pending[0]) if self.pending is not None else (None, (0, 0.0))
        if self.p2p:
            # rows of the previous evaluation sit in the array written then; this evaluation writes the other one
            rows_in = self._rows[self._flip ^ 1] if self.pending is not None else None
            e.bh_step_groups(self.pos, self.vel, self.theta, self.eps, dt, self.g0, self.g1 - self.g0, pre_drift=pre, kick=kick,
                             drift=drift, rows_in=rows_in, prev_drift=prev, rows_peers=[b.ptr for b in self._peers[self._flip]],
                             stats_out=self.stats_buf if stats else None, flags=self.flags, sync=self._sync)
            self.force_evaluations += 1
            self._exchange()                      # barrier: every rank's stores of this evaluation have landed
            self.rows_all = self._rows[self._flip]
            self._flip ^= 1
            self.pending = (drift, dt)
            return
        e.bh_step_groups(self.pos, self.vel, self.theta, self.eps, dt, self.g0, self.g1 - self.g0, pre_drift=pre, kick=kick,
                         drift=drift, rows_in=rows_in, prev_drift=prev, rows_out=self.rows_out if sharded else None,
                         stats_out=self.stats_buf if stats else None, flags=self.flags)
        self.force_evaluations += 1
        if sharded:
            self._exchange()
            self.pending = (drift, dt)

    def materialize(self):
        """Brings pos / vel (body order) up to date on this rank; a no-op on one GPU."""
        if self.pending is not None:
            extra = {}
            if self.p2p and self._sync is not None:     # the scatter waits for every rank's rows on the device
                extra = dict(rows_peers=[b.ptr for b in self._peers[self._flip]], sync=self._sync)
            self.engine.bh_step_groups(self.pos, self.vel, self.theta, self.eps, self.pending[1], 0, 0, rows_in=self.rows_all,
                                       prev_drift=self.pending[0], flags=self.flags, **extra)
            self.pending = None

    def step(self, dt, lazy=False, stats=False):
        pre, groups = stepper_program(self.integrator, dt)
        for k, (kick, drift) in enumerate(groups):
            self._group(dt, pre if k == 0 else (DRIFT_NONE, 0.0), kick, drift, stats and k == len(groups) - 1)
        if not lazy:
            self.materialize()
        if stats:
            return self.kinetic_stats()

    def kinetic_stats(self):
        """(sum m|v|, kinetic energy) of the whole system from the last step(stats=True): per-group partial sums of the
        walk epilogue, added in group order; sharded runs all-reduce the two doubles."""
        import torch
        part = self.stats_buf.view(-1, 2)[self.g0:self.g1].sum(dim=0)
        if self.world > 1:
            self.dist.all_reduce(part)
        m, k = part.tolist()
        return m, k


class FusedDirect:
    """Direct-sum time stepping with the kick and drift fused into the kernels of the sum (nb_direct_step_slice): 3 launches
    per force evaluation instead of 5, replayed as one CUDA graph at launch-bound sizes.  Sharded (world > 1): the rank owns
    the body slice [begin, end); a pre-force drift runs as a kernel of its own so that the positions can be all-gathered
    before the force (the exchange of ShardedSystem, one collective per force evaluation)."""

    def __init__(self, engine, pos, vel, integrator="Verlet_DKD", dim=3, softening=1.01, world=1, rank=0, dist=None, flags=0,
                 stats=False):
        import torch
        from . import shard
        self.engine, self.pos, self.vel = engine, pos, vel
        self.n = pos.shape[0]
        self.integrator, self.dim, self.eps, self.flags = integrator, dim, softening, flags
        self.world, self.rank, self.dist = world, rank, dist
        self.begin, self.end = shard.slice_bounds(self.n, world, rank)
        self.count = self.end - self.begin
        self._shard = shard
        self.stats_buf = torch.zeros(2 * ((self.count + 255) // 256), dtype=torch.float64, device=pos.device) if stats else None
        self.force_evaluations = 0

    def step(self, dt, stats=False):
        pre, groups = stepper_program(self.integrator, dt)
        e = self.engine
        for k, (kick, drift) in enumerate(groups):
            p = pre if k == 0 else (DRIFT_NONE, 0.0)
            if self.world > 1:
                if p[0] != DRIFT_NONE:
                    e.integrate_substep(self.pos, self.vel, self.vel, self.begin, self.count, dt, 0, 0.0, p[0], p[1])
                    p = (DRIFT_NONE, 0.0)
                self._shard.gather_positions(self.pos, self.n, self.world, self.rank, self.dist)
            e.direct_step_slice(self.pos, self.vel, self.dim, self.eps, dt, self.begin, self.count, pre_drift=p, kick=kick,
                                drift=drift, stats_out=self.stats_buf if (stats and k == len(groups) - 1) else None,
                                flags=self.flags)
            self.force_evaluations += 1
        if stats:
            part = self.stats_buf.view(-1, 2).sum(dim=0)
            if self.world > 1:
                self.dist.all_reduce(part)
            return tuple(part.tolist())

    def materialize(self):
        """Replicates the positions of all slices (the last drift of a sharded step is slice-local)."""
        if self.world > 1:
            self._shard.gather_positions(self.pos, self.n, self.world, self.rank, self.dist)
