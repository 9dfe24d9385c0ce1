"""Runs a few Barnes-Hut evaluations at N = 1,048,576, theta = 0.5 (for ncu launch lists / captures)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
eng = pkg.Engine(0, n)
pos = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1)).cuda()
acc = torch.zeros(n, 4, device="cuda")
for _ in range(reps):
    eng.bh_accel(pos, acc, 0.5, 1.01)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(reps):
    eng.bh_accel(pos, acc, 0.5, 1.01)
b.record()
torch.cuda.synchronize()
print("ms per evaluation:", a.elapsed_time(b) / reps)
