"""torchrun --nproc-per-node G profiles/allgather_probe.py : what the collectives a sharded tree build would need cost on this
box: in-place all-gathers of 16 MB (positions or rows at N = 1M), 52 MB (all node records at N = 1M),
268 MB (rows at N = 16.7M) and 830 MB (node records at N = 16.7M), and a 4-byte all-reduce (the barrier behind the peer
stores).  CUDA events on the launching stream, max over ranks, 20 repetitions after 5 warm-ups."""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = {"n_gpus": world}
for name, total in (("positions_1m_16MB", 16 << 20), ("tree_1m_52MB", 52 << 20), ("rows_16m_268MB", 268 << 20),
                    ("tree_16m_830MB", 830 << 20)):
    per = total // world // 16 * 16
    buf = torch.empty(per * world, dtype=torch.uint8, device="cuda")
    mine = buf[rank * per:(rank + 1) * per]
    for _ in range(5):
        dist.all_gather_into_tensor(buf, mine)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        dist.all_gather_into_tensor(buf, mine)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 20], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    out[name] = {"ms": ms, "received_GBps_per_rank": per * (world - 1) / (ms * 1e-3) / 1e9}
    del buf
tok = torch.zeros(1, device="cuda")
for _ in range(5):
    dist.all_reduce(tok)
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    dist.all_reduce(tok)
e1.record()
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / 50], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
out["all_reduce_4B_back_to_back_ms"] = float(t.item())
if rank == 0:
    print("PROBE " + json.dumps(out), flush=True)
dist.destroy_process_group()
