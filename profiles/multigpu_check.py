"""torchrun --nproc-per-node G profiles/multigpu_check.py : sharded force evaluations against the single-GPU
result on rank 0 (direct ordered mode: bit-identical; Barnes-Hut: bit-identical, the tree is rebuilt
redundantly and the walk is deterministic)."""
import os
import sys

# Small systems let 2 or 4 warps share a group (NB_BH_WALK_SPLIT, chosen from the number of groups a GPU walks), which
# changes the ORDER of a target's FP32 sum with the number of ranks.  One warp per group everywhere makes the sharded
# result comparable bit for bit; the last block below compares the default schedule within rounding.
os.environ["NB_BH_WALK_SPLIT"] = "1"

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = pkg.Engine(local, 0)
ok = True
for n in (100000, 65536):
    pos_h = pkg.ics.uniform_disk(n, seed=9, z_half_height=3000.0)
    b, e = pkg.shard.slice_bounds(n, world, rank)
    # every rank starts with only ITS slice valid; the all-gather must fill in the rest
    pos = torch.zeros(n, 4, device="cuda")
    pos[b:e] = torch.from_numpy(pos_h[b:e]).cuda()
    pkg.shard.gather_positions(pos, n, world, rank, dist)
    assert torch.equal(pos.cpu(), torch.from_numpy(pos_h)), "all-gather did not replicate the positions"
    acc = torch.zeros(e - b, 4, device="cuda")
    eng.direct_accel_slice(pos, b, e - b, acc, 3, 1.01, pkg.NB_DIRECT_ORDERED)
    full = torch.zeros(n, 4, device="cuda")
    eng.direct_accel(pos, full, 3, 1.01, pkg.NB_DIRECT_ORDERED)
    ok &= bool(torch.equal(acc, full[b:e]))
    # Barnes-Hut: sharded walk, then sum-reduce the (disjoint) contributions to compare with the full walk
    pos2 = pos.clone()
    pos2[:, 2] = 0
    part = torch.zeros(n, 4, device="cuda")
    g0, g1 = pkg.shard.group_bounds(n, world, rank)
    eng.bh_accel_groups(pos2, part, 0.5, 1.01, g0, g1 - g0)
    dist.all_reduce(part)
    whole = torch.zeros(n, 4, device="cuda")
    eng.bh_accel(pos2, whole, 0.5, 1.01)
    ok &= bool(torch.equal(part, whole))
# fused sharded stepping (nb_bh_step_groups + one all-gather of 16 B/body rows per force evaluation): after every step
# each rank's replicated state equals the one-GPU fused trajectory bit for bit, for even and ragged group counts
import numpy as np  # noqa: E402
for n, name, steps in ((65536, "Verlet_DKD", 3), (100001, "PEFRL", 2), (50000, "ForestRuth", 2)):
    pos_h = pkg.ics.uniform_disk(n, seed=21)
    rng = np.random.default_rng(4)
    vel_h = np.zeros((n, 4), np.float32)
    vel_h[:, :2] = rng.standard_normal((n, 2)).astype(np.float32) * 2.0e4
    p1, v1 = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    one = pkg.integrators.FusedBarnesHut(pkg.Engine(local, n), p1, v1, name, 0.5, 1.01)
    pg, vg = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    many = pkg.integrators.FusedBarnesHut(pkg.Engine(local, n), pg, vg, name, 0.5, 1.01, world, rank, dist, stats=True)
    pp, vp = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    peer = pkg.integrators.FusedBarnesHut(pkg.Engine(local, n), pp, vp, name, 0.5, 1.01, world, rank, dist, p2p=True)
    for k in range(steps):
        one.step(1.0e-3)
        many.step(1.0e-3, lazy=(k + 1 < steps))
        peer.step(1.0e-3, lazy=(k + 1 < steps))
    torch.cuda.synchronize()
    same = bool(torch.equal(p1, pg) and torch.equal(v1, vg) and torch.equal(p1, pp) and torch.equal(v1, vp))
    if not same and rank == 0:
        print("fused sharded trajectory differs:", n, name)
    ok &= same
# the fused direct step sharded by body slices (drift kernel, all-gather of positions, force + kick + drift in the sum's own
# kernels): in NB_DIRECT_ORDERED mode every slice reproduces the one-GPU trajectory bit for bit
for n, name, dim in ((20000, "Verlet_DKD", 3), (8192, "PEFRL", 2)):
    pos_h = pkg.ics.uniform_disk(n, seed=23, z_half_height=500.0 if dim == 3 else 0.0)
    rng = np.random.default_rng(6)
    vel_h = np.zeros((n, 4), np.float32)
    vel_h[:, :dim] = rng.standard_normal((n, dim)).astype(np.float32) * 2.0e4
    p1, v1 = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    one = pkg.integrators.FusedDirect(pkg.Engine(local, n), p1, v1, name, dim, 1.01, flags=pkg.NB_DIRECT_ORDERED)
    b, e = pkg.shard.slice_bounds(n, world, rank)
    pg, vg = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h[b:e]).cuda()
    many = pkg.integrators.FusedDirect(pkg.Engine(local, n), pg, vg, name, dim, 1.01, world, rank, dist, flags=pkg.NB_DIRECT_ORDERED)
    for _ in range(2):
        one.step(1.0e-3)
        many.step(1.0e-3)
    many.materialize()
    torch.cuda.synchronize()
    same = bool(torch.equal(p1, pg) and torch.equal(v1[b:e], vg))
    if not same and rank == 0:
        print("fused sharded direct trajectory differs:", n, name)
    ok &= same
t = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("MULTIGPU_CHECK", "PASS" if int(t.item()) == 1 else "FAIL", "world", world)
dist.destroy_process_group()
