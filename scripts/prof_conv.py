"""One resident convolution / one loop step on a volume that tiles exactly, for ncu captures.
usage: python scripts/prof_conv.py [X Y Z K reps]   (run under gpurun)"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import spim_registration_b200 as m  # noqa: E402
from spim_registration_b200 import _lib as L, synthetic  # noqa: E402
from spim_registration_b200.parallel import ShardedDeconvolution  # noqa: E402

a = [int(x) for x in sys.argv[1:]]
X, Y, Z, K, reps = (a + [480, 480, 452, 31, 2][len(a):])[:5]
lib = m.load()
torch.cuda.set_device(0)
V = 2
if os.environ.get("MVD_PROF_SEP", "0") == "1":  # axis-aligned Gaussians: rank-1 kernels, direct separable path
    def _g(n, sigma):
        x = np.arange(n, dtype=np.float64) - n // 2
        return np.exp(-x * x / (2 * sigma * sigma))
    psfs = [(_g(K, K / 7.0 + 0.3 * v)[:, None, None] * _g(K, K / 8.0)[None, :, None] * _g(K, K / 9.0 + 0.2 * v)[None, None, :]).astype(np.float32)
            for v in range(V)]
else:
    psfs = synthetic.make_psfs(V, (K, K, K))
k1 = [(p / p.sum()).astype(np.float32) for p in psfs]
k2 = [np.ascontiguousarray(k[::-1, ::-1, ::-1]) for k in k1]
g = torch.Generator(device="cuda")
g.manual_seed(1)
imgs = [torch.rand((Z, Y, X), generator=g, device="cuda") * 0.9 + 0.05 for _ in range(V)]
w = [torch.full((Z, Y, X), 1.0 / V, device="cuda") for _ in range(V)]
dec = ShardedDeconvolution((X, Y, Z), imgs, w, k1, k2, 0.006, None, 0)
L.check(lib.mvdSetDirectPath(dec.ctx, int(os.environ.get("MVD_PROF_DIRECT", os.environ.get("MVD_PROF_SEP", "0")))), "mvdSetDirectPath")
a_, b_ = C.c_int(), C.c_int()
L.check(lib.mvdViewUsesDirectPath(dec.ctx, 0, C.byref(a_), C.byref(b_)), "mvdViewUsesDirectPath")
print("view 0 direct-path classes (conv1, conv2):", a_.value, b_.value)
dec.init_psi_average()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
dec.run_iterations(1)
torch.cuda.synchronize()
e0.record()
dec.run_iterations(reps)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print("volume %dx%dx%d K=%d V=%d: %.3f ms/iteration, %.3f Gvoxel-it/s, %.3f ms per convolution" %
      (X, Y, Z, K, V, ms, X * Y * Z / ms / 1e6, ms / (2 * V)))
if os.environ.get("MVD_PROF_MODEL", "1") == "1":
    # byte model of one iteration (2 V convolutions), per kernel: what `roofline.achieved` divides by
    L.check(lib.mvdProfileStart(dec.ctx), "start")
    dec.run_iterations(1)
    torch.cuda.synchronize()
    pms, pl, pb = (C.c_double * 3)(), (C.c_longlong * 3)(), (C.c_double * 3)()
    L.check(lib.mvdProfileStop(dec.ctx, pms, pl, pb), "stop")
    for i, n in enumerate(("fwd_xy", "z_pass", "inv_xy")):
        print("model %s: %d launches, %.3f ms, %.6f GB algorithmic per iteration (%d convolutions)" % (n, pl[i], pms[i], pb[i] / 1e9, 2 * V))
dec.close()
