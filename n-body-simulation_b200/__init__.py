"""B200-native gravitational force engine -- Python host side.

Mirrors the reference's compute plug-in surface (MoonArchitect/N-body-simulation,
src/ComputeMethods.h:8-42, src/NbodySystem.h:12-43) on top of the C ABI in include/nbody_b200.h.
All arithmetic happens in libnbody_b200.so (hand-written sm_100a CUDA); PyTorch is used only for
device memory, streams and torch.distributed.  There is no CPU fallback: without the built library or
without a B200 every compute call raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# NB_LIB: another build of the same library (profiles/build_variant.sh: tuning experiments side by side in one process tree)
_LIB_PATH = os.environ.get("NB_LIB") or os.path.join(_HERE, "libnbody_b200.so")
_lib = None

NB_OK = 0
NB_DIRECT_AUTO, NB_DIRECT_ORDERED, NB_DIRECT_NO_HOIST = 0, 1, 2
NB_BH_FRESH_BOUNDS, NB_BH_ACCUMULATE_BOUNDS, NB_BH_COUNT_WALK, NB_BH_PLANAR = 0, 1, 0x100, 0x200

# every symbol include/nbody_b200.h declares (tests check that the library exports each one)
ABI_SYMBOLS = [
    "nb_create", "nb_destroy", "nb_last_error", "nb_device_info",
    "nb_direct_accel", "nb_direct_accel_slice", "nb_direct_accel_host", "nb_direct_accel_noalloc",
    "nb_direct_last_launch", "nb_direct_step_slice",
    "nb_bh_accel", "nb_bh_accel_groups", "nb_bh_accel_host", "nb_bh_stages_host", "nb_bh_export_tree",
    "nb_bh_last_counters", "nb_bh_group_work", "nb_bh_build_info", "nb_bh_step_groups", "nb_peer_alloc", "nb_peer_open", "nb_peer_close",
    "nb_peer_free", "nb_calibrate", "nb_verlet_dkd_drift", "nb_verlet_dkd_kick_drift",
    "nb_integrate_substep", "nb_stats", "nb_ic_uniform_ellipsoid", "nb_ic_uniform_box", "nb_ic_disk_velocities", "nb_ic_standard", "nb_ic_standard_velocities",
    "nb_bh3d_accel", "nb_bh3d_accel_groups", "nb_bh3d_stages_host", "nb_bh3d_export_tree", "nb_bh3d_last_counters",
]


class NbodyError(RuntimeError):
    pass


class DirectStep(C.Structure):
    """struct nb_direct_step of include/nbody_b200.h."""
    _fields_ = [("softening", C.c_float), ("dt", C.c_float), ("dim", C.c_int), ("flags", C.c_int),
                ("target_begin", C.c_int), ("target_count", C.c_int),
                ("pre_drift_kind", C.c_int), ("pre_drift_coef", C.c_double),
                ("kick_kind", C.c_int), ("kick_coef", C.c_double),
                ("drift_kind", C.c_int), ("drift_coef", C.c_double),
                ("acc_out", C.c_void_p), ("stats_out", C.c_void_p)]


class BHStep(C.Structure):
    """struct nb_bh_step of include/nbody_b200.h."""
    _fields_ = [("theta", C.c_float), ("softening", C.c_float), ("dt", C.c_float), ("flags", C.c_int),
                ("group_begin", C.c_int), ("group_count", C.c_int),
                ("pre_drift_kind", C.c_int), ("pre_drift_coef", C.c_double),
                ("kick_kind", C.c_int), ("kick_coef", C.c_double),
                ("drift_kind", C.c_int), ("drift_coef", C.c_double),
                ("rows_in", C.c_void_p), ("prev_drift_kind", C.c_int), ("prev_drift_coef", C.c_double),
                ("rows_out", C.c_void_p), ("acc_out", C.c_void_p), ("stats_out", C.c_void_p),
                ("n_rows_peers", C.c_int), ("rows_peers", C.c_void_p * 8),
                ("sync_rank", C.c_int), ("sync_local", C.c_void_p), ("sync_peers", C.c_void_p * 8)]


class PeerBuffer:
    """Device memory other ranks can map (nb_peer_alloc / nb_peer_open).  `tensor` is a torch view of it (no copy)."""

    def __init__(self, engine, shape, ptr, handle=None, opened=False):
        self.engine, self.shape, self.ptr, self.handle, self.opened = engine, tuple(shape), ptr, handle, opened
        n = 1
        for d in self.shape:
            n *= d
        self.__cuda_array_interface__ = {"shape": self.shape, "typestr": "<f4", "data": (ptr, False), "version": 2,
                                         "strides": None}
        self.bytes = 4 * n

    @property
    def tensor(self):
        import torch
        return torch.as_tensor(self, device=torch.device("cuda", self.engine.device))

    def release(self):
        if self.ptr:
            (lib().nb_peer_close if self.opened else lib().nb_peer_free)(self.engine._h, C.c_void_p(self.ptr))
            self.ptr = 0


def library_path():
    return _LIB_PATH


def lib():
    """Loads the C-ABI library.  Raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise NbodyError(f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(the engine has no CPU or PyTorch fallback)")
        L = C.CDLL(_LIB_PATH)
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.nb_create.argtypes = [C.POINTER(vp), i, i]
        L.nb_destroy.argtypes = [vp]
        L.nb_last_error.argtypes = [vp]
        L.nb_last_error.restype = C.c_char_p
        L.nb_device_info.argtypes = [vp, C.POINTER(i), C.POINTER(i), C.POINTER(C.c_uint64)]
        L.nb_direct_accel.argtypes = [vp, vp, vp, i, i, f, i, vp]
        L.nb_direct_accel_slice.argtypes = [vp, vp, i, i, i, vp, i, f, i, vp]
        L.nb_direct_accel_host.argtypes = [vp, vp, vp, i, i, f, i]
        L.nb_direct_last_launch.argtypes = [vp, C.POINTER(i), C.POINTER(i), C.POINTER(i)]
        L.nb_direct_step_slice.argtypes = [vp, vp, vp, i, C.POINTER(DirectStep), vp]
        L.nb_bh_accel.argtypes = [vp, vp, vp, i, f, f, i, vp]
        L.nb_bh_accel_groups.argtypes = [vp, vp, vp, i, f, f, i, i, i, vp]
        L.nb_bh_accel_host.argtypes = [vp, vp, vp, i, f, f, i]
        L.nb_bh_stages_host.argtypes = [vp, vp, i, i, vp, vp, vp, vp]
        L.nb_bh_export_tree.argtypes = [vp, C.POINTER(i), vp, vp, vp, vp, vp, vp]
        L.nb_bh_last_counters.argtypes = [vp, C.POINTER(C.c_int64)]
        L.nb_bh_build_info.argtypes = [vp, C.POINTER(C.c_int64)]
        L.nb_bh_group_work.argtypes = [vp, vp, i]
        L.nb_bh_step_groups.argtypes = [vp, vp, vp, i, C.POINTER(BHStep), vp]
        L.nb_peer_alloc.argtypes = [vp, C.c_uint64, C.POINTER(vp), C.c_char_p]
        L.nb_peer_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
        L.nb_peer_close.argtypes = [vp, vp]
        L.nb_peer_free.argtypes = [vp, vp]
        L.nb_bh3d_accel.argtypes = [vp, vp, vp, i, f, f, i, vp]
        L.nb_bh3d_accel_groups.argtypes = [vp, vp, vp, i, f, f, i, i, i, vp]
        L.nb_bh3d_stages_host.argtypes = [vp, vp, i, vp, vp, vp]
        L.nb_bh3d_export_tree.argtypes = [vp, C.POINTER(i), vp, vp, vp, vp, vp, vp, vp]
        L.nb_bh3d_last_counters.argtypes = [vp, C.POINTER(C.c_int64)]
        L.nb_calibrate.argtypes = [vp, C.POINTER(C.c_double)]
        L.nb_verlet_dkd_drift.argtypes = [vp, vp, vp, i, i, f, vp]
        L.nb_verlet_dkd_kick_drift.argtypes = [vp, vp, vp, vp, i, i, f, vp]
        d = C.c_double
        L.nb_ic_uniform_ellipsoid.argtypes = [vp, vp, vp, i, i, C.c_uint64, d, d, d, d, d, f, f, f, f, vp]
        L.nb_ic_uniform_box.argtypes = [vp, vp, vp, i, i, C.c_uint64, d, d, d, d, i, d, d, d, d, f, f, vp]
        L.nb_ic_disk_velocities.argtypes = [vp, vp, vp, vp, i, i, f, f, i, f, f, vp]
        L.nb_ic_standard.argtypes = [vp, vp, i, i, C.c_uint64, d, d, vp]
        L.nb_ic_standard_velocities.argtypes = [vp, vp, vp, vp, i, i, C.c_uint64, f, vp]
        L.nb_stats.argtypes = [vp, vp, vp, i, i, i, i, C.POINTER(C.c_double), vp]
        L.nb_integrate_substep.argtypes = [vp, vp, vp, vp, vp, i, i, f, i, C.c_double, i, C.c_double, vp]
        _lib = L
    return _lib


def _stream_ptr(stream):
    if stream is None:
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


class Engine:
    """Owner of one nb_context (scratch memory of one GPU)."""

    def __init__(self, device=0, max_bodies=0):
        self._h = C.c_void_p()
        rc = lib().nb_create(C.byref(self._h), int(device), int(max_bodies))
        if rc != NB_OK:
            msg = lib().nb_last_error(None).decode()
            self._h = None
            raise NbodyError(f"nb_create failed ({rc}): {msg}")
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            lib().nb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != NB_OK:
            raise NbodyError(f"{what} failed ({rc}): {lib().nb_last_error(self._h).decode()}")

    # -- direct ------------------------------------------------------------------------------------
    def direct_accel(self, pos, acc, dim, softening, flags=NB_DIRECT_AUTO, stream=None):
        """pos, acc: CUDA float32 tensors [n,4] (x,y,z,m) / (ax,ay,az,0)."""
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(acc, n)
        self._check(lib().nb_direct_accel(self._h, pos.data_ptr(), acc.data_ptr(), n, dim, softening, flags,
                                          _stream_ptr(stream)), "nb_direct_accel")

    def direct_accel_slice(self, pos, target_begin, target_count, acc, dim, softening, flags=NB_DIRECT_AUTO,
                           stream=None):
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(acc, target_count)
        self._check(lib().nb_direct_accel_slice(self._h, pos.data_ptr(), n, target_begin, target_count,
                                                acc.data_ptr(), dim, softening, flags, _stream_ptr(stream)),
                    "nb_direct_accel_slice")

    def direct_accel_host(self, pos, dim, softening, flags=NB_DIRECT_AUTO, out=None):
        """pos: host float32 array/tensor [n,4]; returns host accelerations [n,4] (H2D + kernel + D2H)."""
        p, n = _host_ptr(pos)
        if out is None:
            out = np.empty((n, 4), np.float32)
        o, _ = _host_ptr(out)
        self._check(lib().nb_direct_accel_host(self._h, p, o, n, dim, softening, flags), "nb_direct_accel_host")
        return out

    def direct_step_slice(self, pos, vel, dim, softening, dt, target_begin, target_count, pre_drift=(0, 0.0), kick=(0, 0.0),
                          drift=(0, 0.0), acc_out=None, stats_out=None, flags=NB_DIRECT_AUTO, stream=None):
        """One [pre-drift; force; kick; drift] group of a stepper with the kick and drift fused into the kernels of the
        direct sum (nb_direct_step_slice).  pos: CUDA float32 [n,4]; vel: [target_count,4], slice-local."""
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(vel, target_count)
        st = DirectStep(softening, dt, dim, flags, target_begin, target_count, pre_drift[0], pre_drift[1], kick[0], kick[1],
                        drift[0], drift[1], acc_out.data_ptr() if acc_out is not None else None,
                        stats_out.data_ptr() if stats_out is not None else None)
        self._check(lib().nb_direct_step_slice(self._h, pos.data_ptr(), vel.data_ptr(), n, C.byref(st), _stream_ptr(stream)),
                    "nb_direct_step_slice")

    def direct_last_launch(self):
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        lib().nb_direct_last_launch(self._h, C.byref(a), C.byref(b), C.byref(c))
        return dict(jsplit=a.value, blocks=b.value, launches=c.value)

    # -- Barnes-Hut ----------------------------------------------------------------------------------
    def bh_accel(self, pos, acc, theta, softening, flags=NB_BH_FRESH_BOUNDS, stream=None):
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(acc, n)
        self._check(lib().nb_bh_accel(self._h, pos.data_ptr(), acc.data_ptr(), n, theta, softening, flags,
                                      _stream_ptr(stream)), "nb_bh_accel")

    def bh_accel_groups(self, pos, acc, theta, softening, group_begin, group_count, flags=NB_BH_FRESH_BOUNDS,
                        stream=None):
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(acc, n)
        self._check(lib().nb_bh_accel_groups(self._h, pos.data_ptr(), acc.data_ptr(), n, theta, softening, flags,
                                             group_begin, group_count, _stream_ptr(stream)), "nb_bh_accel_groups")

    def bh_accel_host(self, pos, theta, softening, flags=NB_BH_FRESH_BOUNDS, out=None):
        p, n = _host_ptr(pos)
        if out is None:
            out = np.empty((n, 4), np.float32)
        o, _ = _host_ptr(out)
        self._check(lib().nb_bh_accel_host(self._h, p, o, n, theta, softening, flags), "nb_bh_accel_host")
        return out

    def bh_stages(self, pos, flags=NB_BH_FRESH_BOUNDS):
        """Runs bounds -> keys -> sorts on device tensor `pos`; returns host arrays."""
        n = pos.shape[0]
        _check_tensor(pos, n)
        bounds = np.zeros(4, np.float32)
        keys = np.zeros(n, np.uint32)
        sidx = np.zeros(n, np.int32)
        tkeys = np.zeros(n, np.uint64)
        self._check(lib().nb_bh_stages_host(self._h, pos.data_ptr(), n, flags, bounds.ctypes.data, keys.ctypes.data,
                                            sidx.ctypes.data, tkeys.ctypes.data), "nb_bh_stages_host")
        return dict(bounds=bounds, keys=keys, sorted_idx=sidx, tree_keys=tkeys)

    def bh_export_tree(self):
        k = C.c_int()
        self._check(lib().nb_bh_export_tree(self._h, C.byref(k), None, None, None, None, None, None),
                    "nb_bh_export_tree")
        k = k.value
        out = dict(prefix=np.zeros(k, np.uint64), level=np.zeros(k, np.int32), count=np.zeros(k, np.int32),
                   mass=np.zeros(k, np.float64), cx=np.zeros(k, np.float64), cy=np.zeros(k, np.float64))
        kk = C.c_int(k)
        self._check(lib().nb_bh_export_tree(self._h, C.byref(kk), out["prefix"].ctypes.data, out["level"].ctypes.data,
                                            out["count"].ctypes.data, out["mass"].ctypes.data, out["cx"].ctypes.data,
                                            out["cy"].ctypes.data), "nb_bh_export_tree")
        return out

    def bh_last_counters(self):
        c = (C.c_int64 * 6)()
        self._check(lib().nb_bh_last_counters(self._h, c), "nb_bh_last_counters")
        return dict(opened=c[0], examined=c[1], interactions=c[2], max_stack=c[3], nodes=c[4], status=c[5])

    def bh_step_groups(self, pos, vel, theta, softening, dt, group_begin, group_count, pre_drift=(0, 0.0), kick=(0, 0.0),
                       drift=(0, 0.0), rows_in=None, prev_drift=(0, 0.0), rows_out=None, acc_out=None, stats_out=None,
                       flags=NB_BH_FRESH_BOUNDS, stream=None, rows_peers=None, sync=None):
        """One [pre-drift; force; kick; drift] group of a stepper with the kick and drift fused into the walk
        (nb_bh_step_groups).  pos, vel: CUDA float32 [n,4], updated in place unless rows_out is given.
        sync = (rank, [device pointer of every rank's sync array]): arrival words instead of a barrier between calls."""
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(vel, n)
        st = BHStep(theta, softening, dt, flags, group_begin, group_count, pre_drift[0], pre_drift[1], kick[0], kick[1],
                    drift[0], drift[1], rows_in.data_ptr() if rows_in is not None else None, prev_drift[0], prev_drift[1],
                    rows_out.data_ptr() if rows_out is not None else None,
                    acc_out.data_ptr() if acc_out is not None else None,
                    stats_out.data_ptr() if stats_out is not None else None)
        if rows_peers:   # device pointers (ints) of every rank's rows array, this rank's own included
            st.n_rows_peers = len(rows_peers)
            for k, ptr in enumerate(rows_peers):
                st.rows_peers[k] = ptr
        if sync is not None:
            st.sync_rank = sync[0]
            st.sync_local = sync[1][sync[0]]
            for k, ptr in enumerate(sync[1]):
                st.sync_peers[k] = ptr
        self._check(lib().nb_bh_step_groups(self._h, pos.data_ptr(), vel.data_ptr(), n, C.byref(st), _stream_ptr(stream)),
                    "nb_bh_step_groups")

    def bh_group_work(self, n):
        """Records summed per target, group by group, in the last NB_BH_COUNT_WALK evaluation (int32 [ceil(n/32)])."""
        w = np.zeros((n + 31) // 32, np.int32)
        self._check(lib().nb_bh_group_work(self._h, w.ctypes.data, w.shape[0]), "nb_bh_group_work")
        return w

    def peer_alloc(self, shape):
        """float32 device buffer of `shape` that other ranks can map; returns a PeerBuffer (handle = 64 bytes to send)."""
        n = 1
        for d in shape:
            n *= d
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        self._check(lib().nb_peer_alloc(self._h, 4 * n, C.byref(ptr), handle), "nb_peer_alloc")
        return PeerBuffer(self, shape, ptr.value, handle.raw)

    def peer_open(self, handle, shape):
        ptr = C.c_void_p()
        self._check(lib().nb_peer_open(self._h, handle, C.byref(ptr)), "nb_peer_open")
        return PeerBuffer(self, shape, ptr.value, handle, opened=True)

    def bh_build_info(self):
        c = (C.c_int64 * 4)()
        self._check(lib().nb_bh_build_info(self._h, c), "nb_bh_build_info")
        return dict(nodes=c[0], repaired=c[1], full_sort_on_device=bool(c[2] & 1), full_sort_by_host=bool(c[2] & 2),
                    deepest_level=c[3])

    # -- Leapfrog (Verlet_DKD) kick/drift on a slice ---------------------------------------------------
    def verlet_dkd_drift(self, pos, vel, begin, count, dt, stream=None):
        self._check(lib().nb_verlet_dkd_drift(self._h, pos.data_ptr(), vel.data_ptr(), begin, count, dt,
                                              _stream_ptr(stream)), "nb_verlet_dkd_drift")

    def verlet_dkd_kick_drift(self, pos, vel, acc, begin, count, dt, stream=None):
        self._check(lib().nb_verlet_dkd_kick_drift(self._h, pos.data_ptr(), vel.data_ptr(), acc.data_ptr(), begin,
                                                   count, dt, _stream_ptr(stream)), "nb_verlet_dkd_kick_drift")

    def integrate_substep(self, pos, vel, acc, begin, count, dt, kick=0, kick_coef=0.0, drift=0, drift_coef=0.0,
                          acc_prev=None, stream=None):
        self._check(lib().nb_integrate_substep(self._h, pos.data_ptr(), vel.data_ptr(), acc.data_ptr(),
                                               acc_prev.data_ptr() if acc_prev is not None else None, begin, count, dt,
                                               kick, kick_coef, drift, drift_coef, _stream_ptr(stream)),
                    "nb_integrate_substep")

    def stats(self, pos, vel, dim, target_begin=0, target_count=None, stream=None):
        """(moment, kinetic energy, potential energy) of a target slice, SimulationStats::computeStats."""
        n = pos.shape[0]
        cnt = n - target_begin if target_count is None else target_count
        out = (C.c_double * 3)()
        self._check(lib().nb_stats(self._h, pos.data_ptr(), vel.data_ptr(), n, target_begin, cnt, dim, out,
                                   _stream_ptr(stream)), "nb_stats")
        return out[0], out[1], out[2]

    # -- device-side initial conditions ---------------------------------------------------------------------
    def ic_uniform_ellipsoid(self, pos, offset, count, seed=1, centre=(0.0, 0.0), radius=(30000.0, 30000.0),
                             z_half_height=0.0, mass=70000.0, central_mass=10000.0, vel=None, bulk_velocity=(0.0, 0.0),
                             stream=None):
        self._check(lib().nb_ic_uniform_ellipsoid(self._h, pos.data_ptr(), vel.data_ptr() if vel is not None else None,
                                                  offset, count, seed, centre[0], centre[1], radius[0], radius[1],
                                                  z_half_height, mass, central_mass, bulk_velocity[0], bulk_velocity[1],
                                                  _stream_ptr(stream)), "nb_ic_uniform_ellipsoid")

    def ic_uniform_box(self, pos, offset, count, seed=1, centre=(0.0, 0.0), size=(60000.0, 60000.0), dim=2,
                       mass_range=(70000.0, 70000.0), speed_range=(0.0, 0.0), bulk_velocity=(0.0, 0.0), vel=None, stream=None):
        """UniformBox (src/InitialConditions.cu:78-107) written straight into device memory."""
        self._check(lib().nb_ic_uniform_box(self._h, pos.data_ptr(), vel.data_ptr() if vel is not None else None, offset, count,
                                            seed, centre[0], centre[1], size[0], size[1], dim, mass_range[0], mass_range[1],
                                            speed_range[0], speed_range[1], bulk_velocity[0], bulk_velocity[1],
                                            _stream_ptr(stream)), "nb_ic_uniform_box")

    def ic_disk_velocities(self, pos, acc, vel, begin, count, centre=(0.0, 0.0), dim=2, bulk_velocity=(0.0, 0.0), stream=None):
        self._check(lib().nb_ic_disk_velocities(self._h, pos.data_ptr(), acc.data_ptr(), vel.data_ptr(), begin, count,
                                                centre[0], centre[1], dim, bulk_velocity[0], bulk_velocity[1],
                                                _stream_ptr(stream)), "nb_ic_disk_velocities")

    def ic_standard(self, pos, offset, count, seed=1, centre=(0.0, 0.0), stream=None):
        """Standard (src/InitialConditions.cu:13-47) written straight into device memory."""
        self._check(lib().nb_ic_standard(self._h, pos.data_ptr(), offset, count, seed, centre[0], centre[1], _stream_ptr(stream)),
                    "nb_ic_standard")

    def ic_standard_velocities(self, pos, acc, vel, offset, count, seed=1, rotation=1.0, stream=None):
        self._check(lib().nb_ic_standard_velocities(self._h, pos.data_ptr(), acc.data_ptr(), vel.data_ptr(), offset, count, seed,
                                                    rotation, _stream_ptr(stream)), "nb_ic_standard_velocities")

    # -- Barnes-Hut in 3D (octree; no reference counterpart) --------------------------------------------------
    def bh3d_accel(self, pos, acc, theta, softening, flags=0, stream=None):
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(acc, n)
        self._check(lib().nb_bh3d_accel(self._h, pos.data_ptr(), acc.data_ptr(), n, theta, softening, flags,
                                        _stream_ptr(stream)), "nb_bh3d_accel")

    def bh3d_accel_groups(self, pos, acc, theta, softening, group_begin, group_count, flags=0, stream=None):
        n = pos.shape[0]
        _check_tensor(pos, n), _check_tensor(acc, n)
        self._check(lib().nb_bh3d_accel_groups(self._h, pos.data_ptr(), acc.data_ptr(), n, theta, softening, flags,
                                               group_begin, group_count, _stream_ptr(stream)), "nb_bh3d_accel_groups")

    def bh3d_stages(self, pos):
        n = pos.shape[0]
        _check_tensor(pos, n)
        bounds = np.zeros(6, np.float32)
        tkeys = np.zeros(n, np.uint64)
        sidx = np.zeros(n, np.int32)
        self._check(lib().nb_bh3d_stages_host(self._h, pos.data_ptr(), n, bounds.ctypes.data, tkeys.ctypes.data,
                                              sidx.ctypes.data), "nb_bh3d_stages_host")
        return dict(bounds=bounds, tree_keys=tkeys, sorted_idx=sidx)

    def bh3d_export_tree(self):
        k = C.c_int()
        self._check(lib().nb_bh3d_export_tree(self._h, C.byref(k), None, None, None, None, None, None, None),
                    "nb_bh3d_export_tree")
        k = k.value
        out = dict(prefix=np.zeros(k, np.uint64), level=np.zeros(k, np.int32), count=np.zeros(k, np.int32),
                   mass=np.zeros(k, np.float64), cx=np.zeros(k, np.float64), cy=np.zeros(k, np.float64),
                   cz=np.zeros(k, np.float64))
        kk = C.c_int(k)
        self._check(lib().nb_bh3d_export_tree(self._h, C.byref(kk), out["prefix"].ctypes.data, out["level"].ctypes.data,
                                              out["count"].ctypes.data, out["mass"].ctypes.data, out["cx"].ctypes.data,
                                              out["cy"].ctypes.data, out["cz"].ctypes.data), "nb_bh3d_export_tree")
        return out

    def bh3d_last_counters(self):
        c = (C.c_int64 * 6)()
        self._check(lib().nb_bh3d_last_counters(self._h, c), "nb_bh3d_last_counters")
        return dict(opened=c[0], examined=c[1], interactions=c[2], max_stack=c[3], nodes=c[4], status=c[5])

    def calibrate(self):
        out = (C.c_double * 4)()
        self._check(lib().nb_calibrate(self._h, out), "nb_calibrate")
        return dict(ffma_tflops=out[0], ffma2_tflops=out[1], mufu_gops=out[2], sm_mhz=out[3])

    def device_info(self):
        d, s, b = C.c_int(), C.c_int(), C.c_uint64()
        lib().nb_device_info(self._h, C.byref(d), C.byref(s), C.byref(b))
        return dict(device=d.value, sm_count=s.value, scratch_bytes=b.value)


def _check_tensor(t, n):
    import torch
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
            and t.dim() == 2 and t.shape[1] == 4 and t.shape[0] >= n):
        raise NbodyError("expected a contiguous CUDA float32 tensor of shape [n, 4]")


def _host_ptr(a):
    try:
        import torch
        if isinstance(a, torch.Tensor):
            if a.is_cuda or a.dtype != torch.float32 or not a.is_contiguous():
                raise NbodyError("expected a contiguous host float32 tensor [n, 4]")
            return C.c_void_p(a.data_ptr()), a.shape[0]
    except ImportError:
        pass
    if not (isinstance(a, np.ndarray) and a.dtype == np.float32 and a.flags.c_contiguous and a.ndim == 2
            and a.shape[1] == 4):
        raise NbodyError("expected a contiguous host float32 array [n, 4]")
    return C.c_void_p(a.ctypes.data), a.shape[0]


from .system import R2, R3, ComputeMethods, NbodySystem  # noqa: E402,F401
from . import ics  # noqa: E402,F401
from . import shard  # noqa: E402,F401
from . import checkpoint  # noqa: E402,F401
from . import integrators  # noqa: E402,F401
from .integrators import IntegrationMethods  # noqa: E402,F401
