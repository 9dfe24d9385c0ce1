import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import __graft_entry__ as g
pkg = g.load_package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, name = 65536, "Verlet_DKD"
pos_h = pkg.ics.uniform_disk(n, seed=21)
rng = np.random.default_rng(4)
vel_h = np.zeros((n, 4), np.float32)
vel_h[:, :2] = rng.standard_normal((n, 2)).astype(np.float32) * 2.0e4
for variant in ("default", "separate_rows", "sync_each"):
    p1, v1 = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    one = pkg.integrators.FusedBarnesHut(pkg.Engine(local, n), p1, v1, name, 0.5, 1.01)
    pg, vg = torch.from_numpy(pos_h).cuda(), torch.from_numpy(vel_h).cuda()
    many = pkg.integrators.FusedBarnesHut(pkg.Engine(local, n), pg, vg, name, 0.5, 1.01, world, rank, dist)
    if variant == "separate_rows":
        many.even = False
        many.rows_out = torch.empty(32 * (many.g1 - many.g0), 4, device="cuda")
    for k in range(3):
        one.step(1.0e-3)
        if variant == "sync_each":
            torch.cuda.synchronize(); dist.barrier()
        many.step(1.0e-3, lazy=False)
        torch.cuda.synchronize()
        dp = (p1 != pg).any(dim=1)
        dv = (v1 != vg).any(dim=1)
        perm = many.engine.bh_stages(pg)["sorted_idx"] if False else None
        print(f"[{variant}] rank {rank} step {k}: pos differ {int(dp.sum())} vel differ {int(dv.sum())} "
              f"max|dp| {float((p1 - pg).abs().max()):.3e} max|dv| {float((v1 - vg).abs().max()):.3e} "
              f"first ids {torch.nonzero(dp)[:5].flatten().tolist()}", flush=True)
    # which rank's groups do the differing bodies belong to?  (sorted position of each body in the last evaluation)
dist.destroy_process_group()
