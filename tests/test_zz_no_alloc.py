"""GPU test of the allocation-free entry point mvdSeparableConvolveNoAlloc (include/mvdecon.h).

The entry point has no counterpart in the reference (added for callers that
cannot call cudaMalloc); it is checked against the existing entry point convolve_N, which runs the same kernels
with a cudaMalloc'ed intermediate.  The file is named so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_l2

pytestmark = pytest.mark.gpu

CONV_TOL = 2e-6   # the tolerance of the single-convolution tests in test_gpu_parity.py
MAX_VOXELS = 8 * 1024 * 1024   # MVD_NOALLOC_MAX_VOXELS


def _kernels(rng, n):
    ks = []
    for _ in range(3):
        k = rng.random(n).astype(np.float32)
        ks.append((k / k.sum()).astype(np.float32))
    return ks


def _no_alloc(lib, L, torch, f, ks, n, oob, oobv):
    """f: numpy [z, y, x]; device buffers come from torch (the caller's side), the library allocates nothing."""
    src = torch.from_numpy(f).cuda()
    out = torch.full_like(src, float("nan"))
    torch.cuda.synchronize()
    kp = [k.ctypes.data_as(L.c_float_p) if k is not None else None for k in ks]
    rc = lib.mvdSeparableConvolveNoAlloc(0, C.c_void_p(src.data_ptr()), C.c_void_p(out.data_ptr()), L.int3(f.shape[::-1]), kp[0], kp[1], kp[2], n,
                                         oob, float(oobv), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    L.check(rc, "mvdSeparableConvolveNoAlloc")
    torch.cuda.synchronize()   # the entry point itself does not synchronise
    return out.cpu().numpy()


def test_no_alloc_separable_convolution_equals_convolve_n_twice_in_a_row(lib):
    torch = pytest.importorskip("torch")
    import spim_registration_b200 as m
    from spim_registration_b200 import _lib as L
    assert lib.getNumDevicesCUDA() >= 1
    cuda = m.CUDASeparableConvolution()
    rng = np.random.default_rng(77)
    # two calls in a row on the shared static scratch: a larger volume first, then a smaller one with other kernels and
    # another out-of-bounds rule, so that anything left over from the first call would show in the second
    cases = [((60, 90, 130), 15, 2, 0.0, True), ((33, 47, 70), 31, 1, 0.6, True), ((20, 40, 50), 7, 0, 0.0, False)]
    for shape, n, oob, oobv, all_axes in cases:
        f = rng.random(shape, dtype=np.float32) + np.float32(0.1)
        ks = _kernels(rng, n)
        if not all_axes:
            ks[1] = None                                   # y axis left alone
        want = f.copy()
        assert cuda.convolve(n, want, ks[0], ks[1], ks[2], shape[2], shape[1], shape[0], True, ks[1] is not None, True, oob, oobv, 0)
        got = _no_alloc(lib, L, torch, f, ks, n, oob, oobv)
        assert np.isfinite(got).all()
        assert rel_l2(got, want) <= CONV_TOL
        np.testing.assert_array_equal(got, want)           # the same kernels on the same data: not a bit differs


def test_no_alloc_rejects_volumes_beyond_the_static_scratch(lib):
    torch = pytest.importorskip("torch")
    from spim_registration_b200 import _lib as L
    k = np.full(7, 1.0 / 7, np.float32)
    kp = k.ctypes.data_as(L.c_float_p)
    dim = (1024, 1024, 9)                                  # 9 Mi voxels > MVD_NOALLOC_MAX_VOXELS
    assert dim[0] * dim[1] * dim[2] > MAX_VOXELS
    src = torch.zeros(16, device="cuda")                   # never touched: the size check comes first
    out = torch.zeros(16, device="cuda")
    rc = lib.mvdSeparableConvolveNoAlloc(0, C.c_void_p(src.data_ptr()), C.c_void_p(out.data_ptr()), L.int3(dim), kp, kp, kp, 7, 0, 0.0, None)
    assert rc != 0
    code, msg = L.last_error()
    assert code != 0 and "exceed the maximum" in msg and str(MAX_VOXELS) in msg
    rc = lib.mvdSeparableConvolveNoAlloc(0, C.c_void_p(src.data_ptr()), C.c_void_p(out.data_ptr()), L.int3((4, 2, 2)), kp, kp, kp, 63, 0, 0.0, None)
    assert rc != 0 and "kernelLen" in L.last_error()[1]
    # a volume of exactly the maximum size is accepted
    n = (256, 256, 128)
    big = torch.rand(n[::-1], device="cuda")
    res = torch.empty_like(big)
    rc = lib.mvdSeparableConvolveNoAlloc(0, C.c_void_p(big.data_ptr()), C.c_void_p(res.data_ptr()), L.int3(n), kp, None, None, 7, 2, 0.0,
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
    L.check(rc, "mvdSeparableConvolveNoAlloc")
    torch.cuda.synchronize()
    assert torch.isfinite(res).all()
