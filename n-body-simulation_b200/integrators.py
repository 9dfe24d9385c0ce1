"""The reference's steppers (namespace IntegrationMethods, src/IntegrationMethods.{cuh,cu}) on the engine.

Same class names and the same sequence of kernels and force evaluations per step; the elementwise kernels are
nb_integrate_substep calls on the rank's target slice, the force evaluation is whatever
`system.force()` does (direct or Barnes-Hut, preceded by the all-gather of positions when sharded).

`system` must provide: engine, pos (replicated [N,4]), vel / acc (slice-local), begin, count, force(), gather().
"""
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

    def __init__(self, engine, pos, vel, method="direct", dim=3, softening=1.01, theta=0.5, world=1, rank=0, dist=None):
        import torch
        from . import shard
        self.engine, self.pos, self.vel = engine, pos, vel
        self.n = pos.shape[0]
        self.world, self.rank, self.dist = world, rank, dist
        self.begin, self.end = shard.slice_bounds(self.n, world, rank)
        self.count = self.end - self.begin
        self.method, self.dim, self.eps, self.theta = method, dim, softening, theta
        self.acc = torch.zeros(self.count, 4, device=pos.device)
        self._acc_full = torch.zeros(self.n, 4, device=pos.device) if method == "bh" else None
        self._shard = shard

    def gather(self):
        self._shard.gather_positions(self.pos, self.n, self.world, self.rank, self.dist)

    def force(self):
        if self.method == "direct":
            self.engine.direct_accel_slice(self.pos, self.begin, self.count, self.acc, self.dim, self.eps)
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
