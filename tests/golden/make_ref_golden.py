"""Generates the golden fixtures of tests/golden/ from the REFERENCE's own CUDA kernels.

Runs on a B200 box (the reference has no CPU path): loads oracle/_ref/libnbody_ref.so -- the unmodified
src/DirectComputeKernels.cu and src/BarnesHutKernels.cu compiled for sm_100 by oracle/Makefile -- drives
it through oracle/ref_shim.cu with harness-owned buffers, and stores inputs + outputs as .npz.
    gpurun -- python tests/golden/make_ref_golden.py gpurun_out/golden
then copy gpurun_out/golden/*.npz into tests/golden/.  The CPU tests pin the oracle to these files.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
oracle = g.load_oracle()
ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libnbody_ref.so"))
vp, i, f = C.c_void_p, C.c_int, C.c_float
ref.ref_direct.argtypes = [vp, vp, i, i, f]
ref.ref_barnes_hut.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, i, f, f]


def ref_direct(pos_h, dim, eps):
    n = pos_h.shape[0]
    pos = torch.from_numpy(pos_h).cuda()
    acc = torch.zeros(n, 4, device="cuda")
    rc = ref.ref_direct(pos.data_ptr(), acc.data_ptr(), n, dim, eps)
    torch.cuda.synchronize()
    assert rc == 0, rc
    return acc.cpu().numpy()


def ref_bh(pos_h, theta, eps, knodes):
    n = pos_h.shape[0]
    m = knodes * n
    pos = torch.zeros(m, 4, device="cuda")
    pos[:n] = torch.from_numpy(pos_h).cuda()
    acc = torch.zeros(n, 4, device="cuda")
    fmax = float(np.finfo(np.float32).max)
    bounds = torch.tensor([fmax, fmax, -fmax, -fmax], device="cuda")  # fresh box
    index = torch.zeros(1, dtype=torch.int32, device="cuda")
    nodes = torch.zeros(4 * m, dtype=torch.int32, device="cuda")
    sidx = torch.zeros(n, dtype=torch.int32, device="cuda")
    keys = torch.zeros(n, dtype=torch.int32, device="cuda")
    rc = ref.ref_barnes_hut(pos.data_ptr(), acc.data_ptr(), bounds.data_ptr(), index.data_ptr(), nodes.data_ptr(),
                            sidx.data_ptr(), keys.data_ptr(), n, m, theta, eps)
    assert rc == 0, rc
    return dict(acc=acc.cpu().numpy(), bounds=bounds.cpu().numpy(), index=int(index.item()),
                nodes=nodes.cpu().numpy().reshape(m, 4), sorted_idx=sidx.cpu().numpy(),
                sorted_keys=keys.cpu().numpy().view(np.uint32), pos_nodes=pos.cpu().numpy())


def canonical_from_ref(nodes, pos_nodes, n):
    """(prefix, level, count, mass, cx, cy) of every reference cell with >= 2 bodies, sorted by
    (prefix, level).  Node ids are dropped (they are nondeterministic)."""
    cells = []
    sys.setrecursionlimit(100000)

    def visit(node, level, prefix):
        cnt = 0
        for k in range(4):
            c = int(nodes[node, k])
            if c == -1:
                continue
            if c < n:
                cnt += 1
            else:
                cnt += visit(c, level + 1, prefix | (k << (62 - 2 * level)))
        if cnt >= 2:
            cells.append((prefix, level, cnt, float(pos_nodes[node, 3]), float(pos_nodes[node, 0]), float(pos_nodes[node, 1])))
        return cnt

    total = visit(n, 0, 0)
    cells.sort(key=lambda t: (t[0], t[1]))
    arr = lambda j, dt: np.array([c[j] for c in cells], dtype=dt)
    return total, dict(prefix=arr(0, np.uint64), level=arr(1, np.int32), count=arr(2, np.int32),
                       mass=arr(3, np.float32), cx=arr(4, np.float32), cy=arr(5, np.float32))


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    eps = 1.01
    # ---- direct ---------------------------------------------------------------------------------
    for name, n, dim in (("direct2d_n2048", 2048, 2), ("direct3d_n2048", 2048, 3), ("direct3d_n16384", 16384, 3)):
        pos = pkg.ics.uniform_disk(n, seed=1, z_half_height=3000.0 if dim == 3 else 0.0)
        acc = ref_direct(pos, dim, eps)
        o32 = oracle.direct_f32(pos, dim, eps)
        o64 = oracle.direct_f64(pos, dim, eps)
        d = dim
        e_or = np.linalg.norm(acc[:, :d] - o32[:, :d], axis=1) / np.linalg.norm(o64[:, :d], axis=1)
        e_64 = np.linalg.norm(acc[:, :d] - o64[:, :d], axis=1) / np.linalg.norm(o64[:, :d], axis=1)
        print(f"{name}: ref-vs-oracle_f32 max {e_or.max():.3e} med {np.median(e_or):.3e}; ref-vs-f64 max {e_64.max():.3e}")
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), pos=pos, acc=acc[:, :d].copy(), eps=np.float32(eps),
                            dim=np.int32(dim))
    # ---- Barnes-Hut -------------------------------------------------------------------------------
    cases = (("bh_disk_n8192", pkg.ics.uniform_disk(8192, seed=1), 0.5, 3),
             ("bh_disk_n16384_t04", pkg.ics.uniform_disk(16384, seed=2), 0.4, 3),
             ("bh_clustered_n8192", pkg.ics.clustered(8192, seed=5, sigma=300.0), 0.7, 8))
    for name, pos, theta, knodes in cases:
        n = pos.shape[0]
        r = ref_bh(pos, theta, eps, knodes)
        total, canon = canonical_from_ref(r["nodes"], r["pos_nodes"], n)
        print(f"{name}: index={r['index']} (pool {knodes * n}), bodies in tree {total}/{n}, cells {len(canon['level'])}, "
              f"max level {canon['level'].max()}")
        # compare with the oracle straight away
        ob = oracle.bh_bounds(pos)
        ok = oracle.bh_keys(pos, ob)
        sk, si = oracle.sort_pairs(ok)
        print("   bounds equal:", np.array_equal(ob, r["bounds"]), " keys equal:", np.array_equal(sk, r["sorted_keys"]),
              " perm equal:", np.array_equal(si, r["sorted_idx"]))
        t = oracle.Tree(pos, ob, si, prebuilt=False)
        oc = t.canonical()
        same = (len(oc["level"]) == len(canon["level"]) and np.array_equal(oc["prefix"], canon["prefix"])
                and np.array_equal(oc["level"], canon["level"]) and np.array_equal(oc["count"], canon["count"]))
        print("   topology equal:", same, len(oc["level"]), len(canon["level"]))
        if same:
            print("   COM max rel diff:", np.abs(oc["cx"] - canon["cx"]).max() / np.abs(canon["cx"]).max(),
                  " mass max rel diff:", (np.abs(oc["mass"] - canon["mass"]) / canon["mass"]).max())
        for pre in (False, True):
            tp = oracle.Tree(pos, ob, si, prebuilt=pre)
            oa, cnt = tp.walk(si, theta, eps)
            dd = np.linalg.norm(oa[:, :2] - r["acc"][:, :2], axis=1) / np.linalg.norm(r["acc"][:, :2], axis=1)
            print(f"   walk prebuilt={pre}: acc rel diff vs ref: med {np.median(dd):.3e} p99 {np.quantile(dd, 0.99):.3e} "
                  f"max {dd.max():.3e}; counters {cnt}")
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), pos=pos, theta=np.float32(theta), eps=np.float32(eps),
                            knodes=np.int32(knodes), bounds=r["bounds"], sorted_keys=r["sorted_keys"],
                            sorted_idx=r["sorted_idx"], acc=r["acc"][:, :2].copy(), n_in_tree=np.int32(total),
                            **{"cell_" + k: v for k, v in canon.items()})


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
