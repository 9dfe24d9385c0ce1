import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import __graft_entry__ as g
pkg = g.load_package()
n = 1 << 20
pos = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1)).cuda()
vel = torch.zeros(n, 4, device="cuda")
eng = pkg.Engine(0, n)
st = pkg.integrators.FusedBarnesHut(eng, pos, vel, "Verlet_DKD", 0.5, 1.01)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(5): st.step(1e-3)
torch.cuda.synchronize()
ts = []
for _ in range(30):
    flush.fill_(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); st.step(1e-3); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
# back to back (no flush): 50 steps
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50): st.step(1e-3)
b.record(); torch.cuda.synchronize()
print("GAB", os.environ.get("NB_BH_GRAPH_MAX_N"), "median_ms", float(np.median(ts)), "back_to_back_ms", a.elapsed_time(b) / 50)
