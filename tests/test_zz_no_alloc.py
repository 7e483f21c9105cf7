"""gsr_dist2_static (GPU): the allocation-free distCUDA2 entry point against the existing gsr_dist2, both called through
the C-ABI with ctypes.  Same kernels, so the bar is the one of test_knn_gpu.py: bit-exact.  The scratch of the new entry
point is one device array shared by all calls and keeps the last call's contents, so the check is made on calls in a row
with different sizes and data.  This file is named to be collected last."""
import ctypes
import os

import numpy as np
import pytest
import torch

import util

pytestmark = pytest.mark.gpu

MAX_POINTS = 262144          # GSR_DIST2_STATIC_MAX_POINTS (include/gsr.h)
GSR_ETOOBIG = -7

ALLOC_FN = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)


class GsrAlloc(ctypes.Structure):
    _fields_ = [("fn", ALLOC_FN), ("ctx", ctypes.c_void_p)]


def _lib():
    lib = ctypes.CDLL(os.path.join(util.ROOT, "structured-gaussian-splatting_b200", "gsr_b200", "libgsr_b200.so"))
    lib.gsr_dist2.restype = ctypes.c_int
    lib.gsr_dist2.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, GsrAlloc, ctypes.c_void_p]
    lib.gsr_dist2_static.restype = ctypes.c_int
    lib.gsr_dist2_static.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.gsr_last_cuda_error.restype = ctypes.c_char_p
    lib.gsr_strerror.restype = ctypes.c_char_p
    return lib


def _points(P, seed):
    g = torch.Generator().manual_seed(seed)
    pts = torch.rand(P, 3, generator=g) * 4 - 2
    if P > 200:
        pts[11] = pts[5]                     # coincident points count as distance 0
        pts[100:120, 1] = -0.5               # a coplanar cluster
    return pts.cuda().contiguous()


def _existing(lib, pts):
    """gsr_dist2 with an allocator callback (a torch tensor made inside the callback, as the binding does)."""
    keep = []

    def alloc(_ctx, nbytes):
        t = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=pts.device)
        keep.append(t)
        return (t.data_ptr() + 255) & ~255

    out = torch.full((pts.shape[0],), float("nan"), device=pts.device)
    rc = lib.gsr_dist2(pts.shape[0], pts.data_ptr(), out.data_ptr(), GsrAlloc(ALLOC_FN(alloc), None),
                       torch.cuda.current_stream().cuda_stream)
    assert rc == 0, (lib.gsr_strerror(rc), lib.gsr_last_cuda_error())
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _static(lib, pts):
    out = torch.full((pts.shape[0],), float("nan"), device=pts.device)
    rc = lib.gsr_dist2_static(pts.shape[0], pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, (lib.gsr_strerror(rc), lib.gsr_last_cuda_error())
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_static_scratch_entry_point_matches_gsr_dist2_on_calls_in_a_row():
    lib = _lib()
    torch.cuda.set_device(0)
    # big first, then smaller ones over the stale contents, then the maximum size
    for P, seed in ((70001, 1), (5000, 2), (5000, 3), (257, 4), (4, 5), (MAX_POINTS, 6), (70001, 1)):
        pts = _points(P, seed)
        want = _existing(lib, pts)
        got = _static(lib, pts)
        assert np.isfinite(got).all()
        assert np.array_equal(got, want), "P = %d: max rel diff %.3e" % (P, np.abs(got / want - 1).max())
        again = _static(lib, pts)            # the same call twice in a row
        assert np.array_equal(again, want)


def test_static_scratch_entry_point_allocates_nothing_and_rejects_large_sizes():
    lib = _lib()
    torch.cuda.set_device(0)
    pts = _points(30000, 7)
    want = _existing(lib, pts)
    _static(lib, pts)                        # first use: the array's address is looked up, kernels are loaded
    out = torch.empty(pts.shape[0], device=pts.device)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    for _ in range(2):
        rc = lib.gsr_dist2_static(pts.shape[0], pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        assert rc == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == stats0
    assert free1 >= free0, "device memory in use grew by %d bytes across two calls" % (free0 - free1)
    assert np.array_equal(out.cpu().numpy(), want)

    # beyond the stated maximum: a clear error, nothing launched, the output untouched
    big = torch.zeros(MAX_POINTS + 1, 3, device="cuda")
    out_big = torch.full((MAX_POINTS + 1,), -1.0, device="cuda")
    rc = lib.gsr_dist2_static(MAX_POINTS + 1, big.data_ptr(), out_big.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == GSR_ETOOBIG
    assert b"GSR_DIST2_STATIC_MAX_POINTS" in lib.gsr_last_cuda_error()
    torch.cuda.synchronize()
    assert bool((out_big == -1.0).all())
    assert lib.gsr_dist2_static(-1, None, None, None) == -1          # GSR_EINVAL
    assert lib.gsr_dist2_static(0, None, None, None) == 0
    # and the entry point still works after the rejected call
    assert np.array_equal(_static(lib, pts), want)
