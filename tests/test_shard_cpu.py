"""World-size-2 (and 3, ragged) gloo tests of the sharding plumbing: slices tile the system, the in-place
all-gather replicates positions, a sharded Leapfrog step equals the single-process step bit for bit, and the
Barnes-Hut group partition covers every group once.  Force arithmetic here is the CPU oracle (allowed in
tests); the GPU kernels are covered by the -m gpu tests."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EPS, DT = 1.01, 0.001


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _np_drift(pos, vel, begin, count, dt):        # DKD_Verlet_1, src/IntegrationMethods.cu:72-79
    p = pos[begin:begin + count, :3].numpy()
    p[:] = (p.astype(np.float64) + vel[:, :3].numpy().astype(np.float64) * (np.float64(np.float32(dt)) * 0.5)).astype(np.float32)


def _np_kick_drift(pos, vel, acc, begin, count, dt):   # DKD_Verlet_2, src/IntegrationMethods.cu:81-93
    v = vel[:, :3].numpy()
    v[:] = (np.float32(dt) * acc[:, :3].numpy() + v).astype(np.float32)  # one rounding differs from FMA by <= 1 ulp; same on all ranks
    _np_drift(pos, vel, begin, count, dt)


def _worker(rank, world, port, n, steps, out):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    pkg, oracle = g.load_package(), g.load_oracle()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pos = torch.from_numpy(pkg.ics.uniform_disk(n, seed=3, z_half_height=3000.0))
    begin, end = pkg.shard.slice_bounds(n, world, rank)
    vel = torch.zeros(end - begin, 4)
    acc = torch.zeros(end - begin, 4)

    def force(p, b, c, a):
        a.copy_(torch.from_numpy(oracle.direct_f32(p.numpy(), 3, EPS, t0=b, nt=c, threads=1)))

    lf = pkg.shard.ShardedLeapfrog(n, world, rank, force, _np_drift, _np_kick_drift, dist)
    for _ in range(steps):
        lf.step(pos, vel, acc, DT)
    pkg.shard.gather_positions(pos, n, world, rank, dist)
    if rank == 0:
        np.save(out, pos.numpy())
    # every rank must hold identical replicated positions
    chk = pos.clone()
    dist.broadcast(chk, 0)
    assert torch.equal(chk, pos)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 512), (3, 500)])
def test_sharded_leapfrog_equals_single_process(tmp_path, world, n):
    steps = 3
    outs = []
    for w in (1, world):
        out = str(tmp_path / f"pos_w{w}.npy")
        mp.spawn(_worker, args=(w, _free_port(), n, steps, out), nprocs=w, join=True)
        outs.append(np.load(out))
    assert np.array_equal(outs[0], outs[1])


def test_partitions_tile(pkg):
    for n in (1, 31, 32, 33, 1000, 1 << 20, (1 << 24) + 5):
        for world in (1, 2, 3, 4, 8):
            cuts = [pkg.shard.slice_bounds(n, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1
            gc = [pkg.shard.group_bounds(n, world, r) for r in range(world)]
            assert gc[0][0] == 0 and gc[-1][1] == (n + 31) // 32
            assert all(a[1] == b[0] for a, b in zip(gc[:-1], gc[1:]))
