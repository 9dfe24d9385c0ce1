"""ctypes front end of the CPU oracle (oracle/nbody_oracle.c).  TEST INFRASTRUCTURE ONLY.

Allowed importers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.
The product package (n-body-simulation_b200/) must never import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "nbody_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "--no-print-directory", so], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)
        u32p, u64p, i32p, i64p = C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int32), C.POINTER(C.c_int64)
        L.oracle_max_threads.restype = C.c_int
        for name in ("oracle_direct_f32", "oracle_direct_f32_fast"):
            getattr(L, name).argtypes = [f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, f32p, C.c_int]
            getattr(L, name).restype = None
        L.oracle_direct_f64.argtypes = [f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, f64p, C.c_int]
        L.oracle_direct_f64.restype = None
        L.oracle_bh_bounds.argtypes = [f32p, C.c_int, f32p, C.c_int]
        L.oracle_bh_keys.argtypes = [f32p, C.c_int, f32p, u32p]
        L.oracle_bh_tree_keys.argtypes = [f32p, C.c_int, f32p, u64p]
        L.oracle_sort_pairs_u32.argtypes = [u32p, C.c_int, u32p, i32p]
        L.oracle_sort_pairs_u64.argtypes = [u64p, C.c_int, u64p, i32p]
        L.oracle_bh_tree_build.argtypes = [f32p, C.c_int, f32p, i32p, C.c_int]
        L.oracle_bh_tree_build.restype = C.c_void_p
        L.oracle_bh_tree_free.argtypes = [C.c_void_p]
        L.oracle_bh_tree_num_nodes.argtypes = [C.c_void_p]
        L.oracle_bh_tree_num_nodes.restype = C.c_int
        L.oracle_bh_tree_canonical.argtypes = [C.c_void_p, u64p, i32p, i32p, f64p, f64p, f64p]
        L.oracle_bh_tree_canonical.restype = C.c_int
        L.oracle_bh_walk.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_float, C.c_float, f32p, i64p, C.c_int]
        L.oracle_bh_accel.argtypes = [f32p, C.c_int, C.c_float, C.c_float, C.c_int, f32p, f32p, u32p, i32p, i64p, C.c_int]
        L.oracle_bh_accel.restype = C.c_int
        L.oracle_bh3_bounds.argtypes = [f32p, C.c_int, f32p]
        L.oracle_bh3_tree_keys.argtypes = [f32p, C.c_int, f32p, u64p]
        L.oracle_bh3_tree_build.argtypes = [f32p, C.c_int, f32p, i32p]
        L.oracle_bh3_tree_build.restype = C.c_void_p
        L.oracle_bh3_tree_free.argtypes = [C.c_void_p]
        L.oracle_bh3_tree_canonical.argtypes = [C.c_void_p, u64p, i32p, i32p, f64p, f64p, f64p, f64p]
        L.oracle_bh3_tree_canonical.restype = C.c_int
        L.oracle_bh3_walk.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_float, C.c_float, f32p, i64p, C.c_int]
        _LIB = L
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _pos(pos):
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    assert pos.ndim == 2 and pos.shape[1] == 4
    return pos


def max_threads():
    return lib().oracle_max_threads()


def direct_f32(pos, dim, eps, t0=0, nt=None, threads=0, fast=False):
    pos = _pos(pos)
    n = pos.shape[0]
    nt = n - t0 if nt is None else nt
    acc = np.zeros((nt, 4), np.float32)
    fn = lib().oracle_direct_f32_fast if fast else lib().oracle_direct_f32
    fn(_p(pos, C.c_float), n, t0, nt, dim, eps, _p(acc, C.c_float), threads)
    return acc


def direct_f64(pos, dim, eps, t0=0, nt=None, threads=0):
    pos = _pos(pos)
    n = pos.shape[0]
    nt = n - t0 if nt is None else nt
    acc = np.zeros((nt, 3), np.float64)
    lib().oracle_direct_f64(_p(pos, C.c_float), n, t0, nt, dim, float(np.float32(eps)), _p(acc, C.c_double), threads)
    return acc


def bh_bounds(pos, prev=None):
    pos = _pos(pos)
    b = np.zeros(4, np.float32) if prev is None else np.array(prev, np.float32)
    lib().oracle_bh_bounds(_p(pos, C.c_float), pos.shape[0], _p(b, C.c_float), 0 if prev is None else 1)
    return b


def bh_keys(pos, bounds):
    pos = _pos(pos)
    b = np.ascontiguousarray(bounds, np.float32)
    k = np.zeros(pos.shape[0], np.uint32)
    lib().oracle_bh_keys(_p(pos, C.c_float), pos.shape[0], _p(b, C.c_float), _p(k, C.c_uint32))
    return k


def bh_tree_keys(pos, bounds):
    pos = _pos(pos)
    b = np.ascontiguousarray(bounds, np.float32)
    k = np.zeros(pos.shape[0], np.uint64)
    lib().oracle_bh_tree_keys(_p(pos, C.c_float), pos.shape[0], _p(b, C.c_float), _p(k, C.c_uint64))
    return k


def sort_pairs(keys):
    keys = np.ascontiguousarray(keys)
    n = keys.shape[0]
    ko = np.zeros_like(keys)
    io = np.zeros(n, np.int32)
    if keys.dtype == np.uint32:
        lib().oracle_sort_pairs_u32(_p(keys, C.c_uint32), n, _p(ko, C.c_uint32), _p(io, C.c_int32))
    elif keys.dtype == np.uint64:
        lib().oracle_sort_pairs_u64(_p(keys, C.c_uint64), n, _p(ko, C.c_uint64), _p(io, C.c_int32))
    else:
        raise TypeError(keys.dtype)
    return ko, io


class Tree:
    """Oracle quadtree (sequential insertion).  Keeps `pos` alive for the C side."""

    def __init__(self, pos, bounds, order=None, prebuilt=False):
        self.pos = _pos(pos)
        self.bounds = np.ascontiguousarray(bounds, np.float32)
        self.n = self.pos.shape[0]
        o = None if order is None else np.ascontiguousarray(order, np.int32)
        self._h = lib().oracle_bh_tree_build(_p(self.pos, C.c_float), self.n, _p(self.bounds, C.c_float),
                                             None if o is None else _p(o, C.c_int32), 1 if prebuilt else 0)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().oracle_bh_tree_free(self._h)
            self._h = None

    @property
    def num_nodes(self):
        return lib().oracle_bh_tree_num_nodes(self._h)

    def canonical(self):
        """dict of arrays sorted by (prefix, level): prefix u64, level, count, mass, cx, cy."""
        k = lib().oracle_bh_tree_canonical(self._h, None, None, None, None, None, None)
        out = dict(prefix=np.zeros(k, np.uint64), level=np.zeros(k, np.int32), count=np.zeros(k, np.int32),
                   mass=np.zeros(k, np.float64), cx=np.zeros(k, np.float64), cy=np.zeros(k, np.float64))
        lib().oracle_bh_tree_canonical(self._h, _p(out["prefix"], C.c_uint64), _p(out["level"], C.c_int32),
                                       _p(out["count"], C.c_int32), _p(out["mass"], C.c_double),
                                       _p(out["cx"], C.c_double), _p(out["cy"], C.c_double))
        return out

    def walk(self, sorted_idx, theta, eps, g0=0, g1=None, threads=0):
        sidx = np.ascontiguousarray(sorted_idx, np.int32)
        groups = (self.n + 31) // 32
        g1 = groups if g1 is None else g1
        acc = np.zeros((self.n, 4), np.float32)
        counters = np.zeros(4, np.int64)
        lib().oracle_bh_walk(self._h, _p(sidx, C.c_int32), g0, g1, theta, eps, _p(acc, C.c_float),
                             _p(counters, C.c_int64), threads)
        return acc, dict(opened=int(counters[0]), examined=int(counters[1]), interactions=int(counters[2]),
                         max_stack=int(counters[3]))


def bh_accel(pos, theta, eps, prebuilt=False, threads=0):
    """Full pipeline.  Returns dict(acc, bounds, keys, sorted_idx, counters, nodes)."""
    pos = _pos(pos)
    n = pos.shape[0]
    acc = np.zeros((n, 4), np.float32)
    bounds = np.zeros(4, np.float32)
    keys = np.zeros(n, np.uint32)
    sidx = np.zeros(n, np.int32)
    counters = np.zeros(4, np.int64)
    k = lib().oracle_bh_accel(_p(pos, C.c_float), n, theta, eps, 1 if prebuilt else 0, _p(acc, C.c_float),
                              _p(bounds, C.c_float), _p(keys, C.c_uint32), _p(sidx, C.c_int32),
                              _p(counters, C.c_int64), threads)
    return dict(acc=acc, bounds=bounds, keys=keys, sorted_idx=sidx, nodes=k,
                counters=dict(opened=int(counters[0]), examined=int(counters[1]),
                              interactions=int(counters[2]), max_stack=int(counters[3])))


# ---- 3D (octree) extension: no reference counterpart, see nbody_oracle.c -------------------------------------

def bh3_bounds(pos):
    pos = _pos(pos)
    b = np.zeros(6, np.float32)
    lib().oracle_bh3_bounds(_p(pos, C.c_float), pos.shape[0], _p(b, C.c_float))
    return b


def bh3_tree_keys(pos, bounds):
    pos = _pos(pos)
    b = np.ascontiguousarray(bounds, np.float32)
    k = np.zeros(pos.shape[0], np.uint64)
    lib().oracle_bh3_tree_keys(_p(pos, C.c_float), pos.shape[0], _p(b, C.c_float), _p(k, C.c_uint64))
    return k


class Tree3:
    """Oracle octree (sequential insertion, 21-level cap)."""

    def __init__(self, pos, bounds, order=None):
        self.pos = _pos(pos)
        self.bounds = np.ascontiguousarray(bounds, np.float32)
        self.n = self.pos.shape[0]
        o = None if order is None else np.ascontiguousarray(order, np.int32)
        self._h = lib().oracle_bh3_tree_build(_p(self.pos, C.c_float), self.n, _p(self.bounds, C.c_float),
                                              None if o is None else _p(o, C.c_int32))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().oracle_bh3_tree_free(self._h)
            self._h = None

    def canonical(self):
        k = lib().oracle_bh3_tree_canonical(self._h, None, None, None, None, None, None, None)
        out = dict(prefix=np.zeros(k, np.uint64), level=np.zeros(k, np.int32), count=np.zeros(k, np.int32),
                   mass=np.zeros(k, np.float64), cx=np.zeros(k, np.float64), cy=np.zeros(k, np.float64),
                   cz=np.zeros(k, np.float64))
        lib().oracle_bh3_tree_canonical(self._h, _p(out["prefix"], C.c_uint64), _p(out["level"], C.c_int32),
                                        _p(out["count"], C.c_int32), _p(out["mass"], C.c_double),
                                        _p(out["cx"], C.c_double), _p(out["cy"], C.c_double), _p(out["cz"], C.c_double))
        return out

    def walk(self, sorted_idx, theta, eps, g0=0, g1=None, threads=0):
        sidx = np.ascontiguousarray(sorted_idx, np.int32)
        groups = (self.n + 31) // 32
        g1 = groups if g1 is None else g1
        acc = np.zeros((self.n, 4), np.float32)
        counters = np.zeros(4, np.int64)
        lib().oracle_bh3_walk(self._h, _p(sidx, C.c_int32), g0, g1, theta, eps, _p(acc, C.c_float),
                              _p(counters, C.c_int64), threads)
        return acc, dict(opened=int(counters[0]), examined=int(counters[1]), interactions=int(counters[2]),
                         max_stack=int(counters[3]))


def bh3_accel(pos, theta, eps, threads=0):
    """bounds -> 63-bit keys -> stable sort -> octree -> walk (groups of 32 in key order)."""
    pos = _pos(pos)
    b = bh3_bounds(pos)
    k = bh3_tree_keys(pos, b)
    sk, si = sort_pairs(k)
    t = Tree3(pos, b, si)
    acc, cnt = t.walk(si, theta, eps, threads=threads)
    return dict(acc=acc, bounds=b, keys=k, sorted_idx=si, counters=cnt, tree=t)
