"""World-size-2 and -3 gloo tests of FusedBarnesHut's host logic (integrators.py): group ownership, the in-place
all-gather of the (x', y', vx', vy') rows (even group counts) and the per-owner broadcasts (ragged), the pending-rows
hand-over from one force evaluation to the next, and materialize().  The engine is a NumPy stand-in with the semantics of
nb_bh_step_groups whose force is the CPU oracle (allowed in tests); the CUDA entry point itself is covered by
tests/test_fused_step_gpu.py.  A sharded run must equal the single-process run bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EPS, THETA, DT = 1.01, 0.5, 0.001


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class OracleStepEngine:
    """nb_bh_step_groups on the host: same argument meaning, NumPy arithmetic (FP64 products rounded once)."""

    def __init__(self, oracle, I):
        self.oracle, self.I, self.perm = oracle, I, None

    def _kick(self, kind, c, dt, v, a):
        I, f = self.I, np.float64
        d = f(np.float32(dt))
        if kind == I.KICK_NONE:
            return v
        if kind == I.KICK_F32:
            return (v.astype(f) + d * a.astype(f)).astype(np.float32)
        if kind == I.KICK_COEF_F64:
            return (v.astype(f) + c * a.astype(f)).astype(np.float32)
        if kind == I.KICK_DTA_F64:
            return (v.astype(f) + (np.float32(dt) * a).astype(np.float32).astype(f) * c).astype(np.float32)
        raise ValueError(kind)

    def _drift(self, kind, c, dt, p, v):
        I, f = self.I, np.float64
        d = f(np.float32(dt))
        if kind == I.DRIFT_NONE:
            return p
        if kind == I.DRIFT_F32:
            return (p.astype(f) + v.astype(f) * d).astype(np.float32)
        if kind == I.DRIFT_HALF_F64:
            return (p.astype(f) + v.astype(f) * (d * 0.5)).astype(np.float32)
        if kind == I.DRIFT_COEF_F64:
            return (p.astype(f) + c * v.astype(f)).astype(np.float32)
        if kind == I.DRIFT_VDT_F64:
            return (p.astype(f) + ((v * np.float32(dt)).astype(np.float32).astype(f) * c) * 0.5).astype(np.float32)
        raise ValueError(kind)

    def bh_step_groups(self, pos, vel, theta, softening, dt, group_begin, group_count, pre_drift=(0, 0.0), kick=(0, 0.0),
                       drift=(0, 0.0), rows_in=None, prev_drift=(0, 0.0), rows_out=None, acc_out=None, stats_out=None,
                       flags=0, stream=None):
        P, V = pos.numpy(), vel.numpy()
        n = P.shape[0]
        if rows_in is not None:
            R = rows_in.numpy()[:n]
            P[self.perm, 0], P[self.perm, 1] = R[:, 0], R[:, 1]
            V[self.perm, 0], V[self.perm, 1] = R[:, 2], R[:, 3]
            P[:, 2] = self._drift(prev_drift[0], prev_drift[1], dt, P[:, 2], V[:, 2])
        P[:, :3] = self._drift(pre_drift[0], pre_drift[1], dt, P[:, :3], V[:, :3])
        if group_count == 0 and kick[0] == 0 and drift[0] == 0:
            return
        r = self.oracle.bh_accel(P, theta, softening, threads=1)
        self.perm = r["sorted_idx"].astype(np.int64)
        s0, s1 = 32 * group_begin, min(32 * (group_begin + group_count), n)
        b = self.perm[s0:s1]
        a = r["acc"][b].astype(np.float32)
        v = V[b].copy()
        p = P[b].copy()
        v[:, :2] = self._kick(kick[0], kick[1], dt, v[:, :2], a[:, :2])
        p[:, :2] = self._drift(drift[0], drift[1], dt, p[:, :2], v[:, :2])
        if rows_out is None:
            p[:, 2] = self._drift(drift[0], drift[1], dt, p[:, 2], v[:, 2])
            P[b], V[b] = p, v
        else:
            rows_out.numpy()[:s1 - s0] = np.stack([p[:, 0], p[:, 1], v[:, 0], v[:, 1]], axis=1)


def _worker(rank, world, port, n, name, steps, out):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    pkg, oracle = g.load_package(), g.load_oracle()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    pos_h = pkg.ics.uniform_disk(n, seed=9)
    pos_h[:, 2] = rng.standard_normal(n).astype(np.float32)
    vel_h = np.zeros((n, 4), np.float32)
    vel_h[:, :3] = rng.standard_normal((n, 3)).astype(np.float32) * 3.0e3
    pos, vel = torch.from_numpy(pos_h), torch.from_numpy(vel_h)
    st = pkg.integrators.FusedBarnesHut(OracleStepEngine(oracle, pkg.integrators), pos, vel, name, THETA, EPS, world, rank, dist)
    for k in range(steps):
        st.step(DT, lazy=(k % 2 == 0))      # lazy steps leave the rows pending for the next evaluation
    st.materialize()
    chk = torch.cat([pos, vel]).clone()
    dist.broadcast(chk, 0)
    assert torch.equal(chk, torch.cat([pos, vel]))          # replicas agree
    if rank == 0:
        np.save(out, torch.cat([pos, vel]).numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,name", [(2, 512, "Verlet_DKD"), (3, 500, "PEFRL"), (2, 1000, "Euler_Symplectic_DK"),
                                          (3, 96, "ForestRuth")])
def test_fused_sharded_equals_single_process(tmp_path, world, n, name):
    outs = []
    for w in (1, world):
        out = str(tmp_path / f"state_w{w}.npy")
        mp.spawn(_worker, args=(w, _free_port(), n, name, 3, out), nprocs=w, join=True)
        outs.append(np.load(out))
    assert np.array_equal(outs[0], outs[1])


def test_stepper_programs_cover_the_fusable_steppers(pkg):
    I = pkg.integrators
    evals = {"Euler_Symplectic_KD": 1, "Euler_Symplectic_DK": 1, "Verlet_DKD": 1, "ForestRuth": 3, "PEFRL": 4}
    for name, k in evals.items():
        pre, groups = I.stepper_program(name, DT)
        assert len(groups) == k == getattr(I.IntegrationMethods, name).force_evaluations
    for name in ("Velocity_Verlet", "Verlet_KDK"):
        with pytest.raises(ValueError):
            I.stepper_program(name, DT)
