"""Walk counters of one Barnes-Hut evaluation (N, theta from argv)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg = g.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
theta = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
kind = sys.argv[3] if len(sys.argv) > 3 else "disk"
eng = pkg.Engine(0, n)
p = pkg.ics.uniform_disk(n, seed=1) if kind == "disk" else pkg.ics.clustered(n, seed=5, sigma=40.0)
pos = torch.from_numpy(p).cuda()
acc = torch.zeros(n, 4, device="cuda")
eng.bh_accel(pos, acc, theta, 1.01, pkg.NB_BH_COUNT_WALK)
print(kind, n, theta, eng.bh_last_counters())
