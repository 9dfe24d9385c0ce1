"""3D (octree) Barnes-Hut at N = 1,048,576 (thick disc), theta = 0.5: time per evaluation and walk counters."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg = g.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
eng = pkg.Engine(0, n)
pos = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1, z_half_height=3000.0)).cuda()
acc = torch.zeros(n, 4, device="cuda")
for _ in range(3):
    eng.bh3d_accel(pos, acc, 0.5, 1.01)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(reps):
    eng.bh3d_accel(pos, acc, 0.5, 1.01)
b.record()
torch.cuda.synchronize()
print("3D BH ms per evaluation:", a.elapsed_time(b) / reps)
eng.bh3d_accel(pos, acc, 0.5, 1.01, pkg.NB_BH_COUNT_WALK)
print(eng.bh3d_last_counters())
full = torch.zeros(n, 4, device="cuda")
torch.cuda.synchronize()
a.record()
eng.direct_accel(pos, full, 3, 1.01)
b.record()
torch.cuda.synchronize()
print("3D direct ms per evaluation:", a.elapsed_time(b))
d = (acc[:, :3] - full[:, :3]).norm(dim=1) / full[:, :3].norm(dim=1)
print("error vs direct sum: median %.3e p99 %.3e max %.3e" % (d.median().item(), d.quantile(0.99).item(), d.max().item()))
