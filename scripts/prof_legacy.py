"""Level 0 (the reference's existing JNA hook): time convolution3DfftCUDAInPlace per block, host arrays in and out,
the way MVDeconFFTThreads.convolve1BlockCUDA calls it (MVDeconFFTThreads.java:52-74).
usage: python scripts/prof_legacy.py [B K calls]   (run under gpurun)"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import spim_registration_b200 as m  # noqa: E402
from spim_registration_b200 import _lib as L  # noqa: E402

a = [int(x) for x in sys.argv[1:]]
B, K, calls = (a + [256, 31, 12][len(a):])[:3]
lib = m.load()
rng = np.random.default_rng(1)
blocks = [rng.random((B, B, B), dtype=np.float32) for _ in range(2)]
kernels = []
for v in range(2):
    k = rng.random((K, K, K), dtype=np.float32)
    kernels.append((k / k.sum()).astype(np.float32))
dim = (C.c_int * 3)(B, B, B)      # (z, y, x) as CUDAFourierConvolution expects
kdim = (C.c_int * 3)(K, K, K)
ref = m.convolve(blocks[0], kernels[0], "zero")
t = []
for i in range(calls):
    blk = blocks[i % 2].copy()
    k = kernels[i % 2]
    t0 = time.perf_counter()
    lib.convolution3DfftCUDAInPlace(blk.ctypes.data_as(L.c_float_p), dim, k.ctypes.data_as(L.c_float_p), kdim, 0)
    t.append(time.perf_counter() - t0)
    if i == 0:
        err = float(np.abs(blk - ref).max())
        print("first call equals mvdConvolve(zero extension): max abs diff %.3e" % err)
code, msg = L.last_error()
print("last error:", code, msg)
t = np.array(t[2:]) * 1e3
eff = (B - K + 1) ** 3
print("block %d^3, kernel %d^3: %.2f ms per call (min %.2f, max %.2f) -> %.1f calls/s, %.3f Geffective-voxel-convolutions/s"
      % (B, K, t.mean(), t.min(), t.max(), 1e3 / t.mean(), eff / t.mean() / 1e6))
