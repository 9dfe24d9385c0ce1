"""Mirror of the reference's plug-in surface for the force path.

    ComputeMethods.ComputeMethod / Direct / BarnesHut   <- src/ComputeMethods.h:8-42, src/ComputeMethods.cpp:8-53
    NbodySystem (N, M, space, host/device buffers, computeAcceleration, updateHost/DeviceData)
                                                        <- src/NbodySystem.h:12-43, src/NbodySystem.cpp:10-25,70-93

Same names, same argument meaning, same ordering rules: NbodySystem's constructor reads compute.knodes
and calls compute.setSystem(self) BEFORE the buffers exist (src/NbodySystem.cpp:12-24), so setSystem
may only look at N, M and space.  Buffers are torch CUDA tensors of shape [M,4] / [N,4] float32, the
float4 layout of the reference (x, y, z, mass).
"""
import numpy as np

R2, R3 = 0, 1  # enum Space, src/NbodySystem.h:12-15


class _Buffers:
    pos_mass = None
    vel = None
    acc = None


class ComputeMethods:
    class ComputeMethod:
        def __init__(self, knodes, softening):
            self.knodes = int(knodes)          # public in the reference (src/ComputeMethods.h:14)
            self.SOFTENING = float(np.float32(softening))
            self.system = None
            self._engine = None

        def setSystem(self, system):
            raise NotImplementedError

        def computeAcc(self):
            raise NotImplementedError

        def _make_engine(self, system):
            from . import Engine
            self._engine = Engine(system.device_index, system.N)

    class Direct(ComputeMethod):
        """src/ComputeMethods.cpp:13-24."""

        def __init__(self, softening, flags=0):
            super().__init__(1, softening)
            self.flags = flags

        def setSystem(self, system):
            self.system = system
            self._make_engine(system)

        def computeAcc(self):
            s = self.system
            dim = 2 if s.space == R2 else 3
            self._engine.direct_accel(s.device.pos_mass, s.device.acc, dim, self.SOFTENING, self.flags)

    class BarnesHut(ComputeMethod):
        """src/ComputeMethods.cpp:29-53.  knodes is accepted and only sizes M (the tree lives in the
        engine's own scratch, pos_mass[N..M) is left alone)."""

        def __init__(self, theta, knodes, softening, flags=0):
            super().__init__(knodes, softening)
            self.theta = float(np.float32(theta))
            self.flags = flags

        def setSystem(self, system):
            self.system = system
            if system.space == R3:
                # src/ComputeMethods.cpp:35-38: same message, same silent downgrade
                print("\n\n\nBarnes-Hut algorithm is not yet available for simulations in 3D space\nSimulating in 2D ...\n")
                system.space = R2
            self._make_engine(system)

        def computeAcc(self):
            s = self.system
            self._engine.bh_accel(s.device.pos_mass, s.device.acc, self.theta, self.SOFTENING, self.flags)


class NbodySystem:
    """The caller of the force path (src/NbodySystem.cpp).  Only what the path needs is mirrored:
    buffers, host<->device copies and computeAcceleration."""

    def __init__(self, nBodies, space, compute, integrator=None, device_index=0):
        import torch
        self.N = int(nBodies)
        self.M = compute.knodes * self.N           # src/NbodySystem.cpp:12
        self.space = space
        self.compute = compute
        self.integrator = integrator
        self.device_index = device_index
        self.host, self.device = _Buffers(), _Buffers()
        compute.setSystem(self)                    # before the buffers exist, src/NbodySystem.cpp:15
        if integrator is not None:
            integrator.setSystem(self)
        dev = torch.device("cuda", device_index)
        self.host.pos_mass = torch.zeros(self.M, 4, dtype=torch.float32).pin_memory()
        self.host.vel = torch.zeros(self.N, 4, dtype=torch.float32).pin_memory()
        self.host.acc = torch.zeros(self.N, 4, dtype=torch.float32).pin_memory()
        self.device.pos_mass = torch.zeros(self.M, 4, dtype=torch.float32, device=dev)
        self.device.vel = torch.zeros(self.N, 4, dtype=torch.float32, device=dev)
        self.device.acc = torch.zeros(self.N, 4, dtype=torch.float32, device=dev)

    def updateDeviceData(self):                    # src/NbodySystem.cpp:83-87
        self.device.pos_mass.copy_(self.host.pos_mass, non_blocking=True)
        self.device.vel.copy_(self.host.vel, non_blocking=True)
        self.device.acc.copy_(self.host.acc, non_blocking=True)

    def updateHostData(self):                      # src/NbodySystem.cpp:76-81
        import torch
        self.host.pos_mass.copy_(self.device.pos_mass, non_blocking=True)
        self.host.vel.copy_(self.device.vel, non_blocking=True)
        self.host.acc.copy_(self.device.acc, non_blocking=True)
        torch.cuda.synchronize(self.device_index)

    def computeAcceleration(self, sync=False):     # src/NbodySystem.cpp:89-93
        import torch
        self.compute.computeAcc()
        if sync:
            torch.cuda.synchronize(self.device_index)
