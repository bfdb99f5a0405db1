"""Throughput of the input-preparation kernels (device-resident buffers): mvdTransformView and mvdNormalizeWeights.
usage: python scripts/prof_inputs.py [X Y Z views]   (run under gpurun)"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import spim_registration_b200 as m  # noqa: E402
from spim_registration_b200 import _lib as L  # noqa: E402

a = [int(x) for x in sys.argv[1:]]
X, Y, Z, V = (a + [1024, 1024, 512, 7][len(a):])[:4]
lib = m.load()
torch.cuda.set_device(0)
src_dim = (X, Y, Z // 2)  # anisotropic stack, z spacing 2
src = torch.rand(src_dim[::-1], device="cuda")
outs, ws = [], []
fb, fr = (C.c_float * 3)(0, 0, 0), (C.c_float * 3)(40, 40, 20)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ms_t = []
for v in range(V):
    ang = np.deg2rad(v * 180.0 / V)
    reg = np.eye(4)
    reg[:3, :3] = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]]) @ np.diag([1.0, 1.0, 2.0])
    c = np.array([X / 2, Y / 2, Z / 2])
    reg[:3, 3] = c - reg[:3, :3] @ np.array([X / 2, Y / 2, Z / 4])
    inv = np.ascontiguousarray(np.linalg.inv(reg)[:3, :4])
    o = torch.empty((Z, Y, X), device="cuda")
    w = torch.empty((Z, Y, X), device="cuda")
    for rep in range(2):
        torch.cuda.synchronize()
        e0.record()
        L.check(lib.mvdTransformView(0, C.c_void_p(src.data_ptr()), L.int3(src_dim), inv.ctypes.data_as(L.c_double_p), L.int3((0, 0, 0)),
                                     L.int3((X, Y, Z)), fb, fr, 1e-4, C.c_void_p(o.data_ptr()), C.c_void_p(w.data_ptr()), 1, 1), "mvdTransformView")
        e1.record()
        torch.cuda.synchronize()
    ms_t.append(e0.elapsed_time(e1))
    outs.append(o)
    ws.append(w)
n = X * Y * Z
print("mvdTransformView %dx%dx%d, %d views: %.2f ms per view (%.1f Gvoxel/s, %.0f GB/s of output)" %
      (X, Y, Z, V, np.mean(ms_t), n / np.mean(ms_t) / 1e6, 8 * n / np.mean(ms_t) / 1e6))
ptrs = (C.c_void_p * V)(*[w.data_ptr() for w in ws])
mn, avg = C.c_int(), C.c_double()
torch.cuda.synchronize()
e0.record()
L.check(lib.mvdNormalizeWeights(0, ptrs, V, n, 1.0, C.byref(mn), C.byref(avg), 1), "mvdNormalizeWeights")
e1.record()
torch.cuda.synchronize()
t = e0.elapsed_time(e1)
print("mvdNormalizeWeights %d views: %.2f ms (%.0f GB/s read+write), min overlap %d, avg overlap %.2f" % (V, t, 8 * V * n / t / 1e6, mn.value, avg.value))
