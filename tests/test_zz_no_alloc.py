"""nb_direct_accel_noalloc: the direct sum for callers that may not allocate device memory (a host whose
time loop has already claimed the GPU's memory, or one that must never hit cudaMalloc's device-wide
synchronisation).  Its partial sums live in a fixed 16 MiB array inside the library instead of the handle's
scratch.  Checked here against nb_direct_accel, through the C ABI, on two calls in a row -- the second call
finds the first call's partials still in the shared array and must not be disturbed by them.

Collected last on purpose (file name): a run that stops at the first failure has then already run every
other test.  Tolerance: the one test_direct_gpu.py uses for the split-source mode (1e-5 of the field norm,
1e-5 median, 5e-5 worst body); when both entry points choose the same split the results are bit-identical."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = 1.01


def _bind(pkg):
    L = pkg.lib()
    L.nb_direct_accel_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p]
    L.nb_direct_accel_noalloc.restype = C.c_int
    return L


def _close(a, b, dim):
    a = a[:, :dim].astype(np.float64)
    b = b[:, :dim].astype(np.float64)
    field = np.linalg.norm(a - b) / np.linalg.norm(b)
    e = np.linalg.norm(a - b, axis=1) / np.linalg.norm(b, axis=1)
    return field < 1e-5 and np.median(e) < 1e-5 and e.max() < 5e-5


def test_noalloc_matches_direct_accel_twice_in_a_row(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    L = _bind(pkg)
    eng = pkg.Engine(0, 0)
    try:
        # two different systems back to back: the second call reuses the array the first one filled
        for n, dim, seed in ((16384, 3, 51), (5000, 2, 52), (16384, 3, 53)):
            pos_h = pkg.ics.uniform_disk(n, seed=seed, z_half_height=3000.0 if dim == 3 else 0.0)
            pos = torch.from_numpy(pos_h).cuda()
            want = torch.zeros(n, 4, device="cuda")
            eng.direct_accel(pos, want, dim, EPS)
            got = torch.full((n, 4), 9.0, device="cuda")
            stream = torch.cuda.current_stream().cuda_stream
            for call in range(2):
                rc = L.nb_direct_accel_noalloc(pos.data_ptr(), got.data_ptr(), n, dim, EPS, stream)
                assert rc == 0, L.nb_last_error(None)
                torch.cuda.synchronize()
                assert _close(got.cpu().numpy(), want.cpu().numpy(), dim), (n, dim, call)
                assert torch.all(got[:, 3] == 0)
                got.fill_(9.0)
    finally:
        eng.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    L = _bind(pkg)
    dummy = torch.zeros(4, 4, device="cuda")
    rc = L.nb_direct_accel_noalloc(dummy.data_ptr(), dummy.data_ptr(), 262144 + 1, 3, EPS, None)
    assert rc == 5                                           # NB_ERR_TOO_LARGE
    assert b"NB_NOALLOC_MAX_BODIES" in L.nb_last_error(None)
    assert L.nb_direct_accel_noalloc(dummy.data_ptr(), dummy.data_ptr(), 4, 5, EPS, None) == 1   # bad dim
