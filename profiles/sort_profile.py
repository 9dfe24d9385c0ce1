"""bounds + keys + both sorts only (nb_bh_stages_host), for ncu launch lists of the sort kernels."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg = g.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
eng = pkg.Engine(0, n)
pos = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1)).cuda()
for _ in range(3):
    eng.bh_stages(pos)
