"""BASELINE.json config 5 sharded: Barnes-Hut N = 16,777,216 two-disc merger, theta = 0.7, Forest-Ruth (3 force
evaluations per step), one process per GPU (torchrun).  Default: the fused step (nb_bh_step_groups: kick + drift in the walk's
epilogue, ONE all-gather of 16 B/body rows per force evaluation).  `classic` as the fourth argument runs round 1's path
(all-gather of positions, separate kick/drift kernels, accelerations returned to their owners by reduce-scatter)."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
classic = len(sys.argv) > 3 and sys.argv[3] == "classic"
p2p = len(sys.argv) > 3 and sys.argv[3] == "p2p"
eng = pkg.Engine(local, n)
pos = torch.from_numpy(pkg.ics.two_galaxies(n)).cuda()
if classic:
    b, e = pkg.shard.slice_bounds(n, world, rank)
    vel = torch.zeros(e - b, 4, device="cuda")
    s = pkg.integrators.ShardedSystem(eng, pos, vel, "bh", 2, 1.01, 0.7, world, rank, dist if world > 1 else None)
    integ = pkg.IntegrationMethods.ForestRuth()
    integ.setSystem(s)
    step = lambda: integ.integrate(0.001)   # noqa: E731
else:
    vel = torch.zeros(n, 4, device="cuda")
    s = pkg.integrators.FusedBarnesHut(eng, pos, vel, "ForestRuth", 0.7, 1.01, world, rank, dist if world > 1 else None, p2p=p2p)
    step = lambda: s.step(0.001, lazy=True)   # noqa: E731
step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
for _ in range(steps):
    step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
dt = (time.perf_counter() - t0) / steps
if classic:
    s.gather()   # the last kick/drift is slice-local: replicate before looking at all positions
else:
    s.materialize()
chk = pos[:, :2].double().sum().item()
if rank == 0:
    print(json.dumps({"config": "5 Barnes-Hut two-galaxy N=%d theta=0.7 Forest-Ruth" % n, "n_gpus": world,
                      "path": "classic" if classic else ("fused, peer stores" if p2p else "fused, all-gather"), "ms_per_step": dt * 1e3,
                      "steps_per_s": 1 / dt, "position_checksum": chk}), flush=True)
if world > 1:
    dist.destroy_process_group()
