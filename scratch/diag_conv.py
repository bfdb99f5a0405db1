# GPU diagnostic: engine convolution vs numpy oracle over sizes / modes (prints, never asserts)
import sys, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import numpy as np
import spim_registration_b200 as m, oracle_np as o
print(m.CUDATools.queryCUDADetails())
rng = np.random.default_rng(0)
def rel(a, b): return float(np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(b))
cases = [((20, 24, 28), (3, 5, 7)), ((32, 32, 32), (15, 15, 15)), ((40, 50, 60), (7, 7, 7)), ((64, 64, 64), (15, 15, 15)),
         ((70, 130, 150), (15, 11, 9)), ((128, 128, 128), (15, 15, 15)), ((96, 200, 260), (31, 31, 31)), ((33, 37, 41), (4, 6, 8)),
         ((300, 40, 40), (9, 9, 9)), ((600, 36, 36), (31, 5, 5))]
for shape, ks in cases:
    f = rng.random(shape, dtype=np.float32) + 0.1
    k = rng.random(ks, dtype=np.float32); k /= k.sum()
    for mode in ("zero", "const", "mirror"):
        t = time.time(); ref = o.convolve_extended(f, k, mode, 1.0); t1 = time.time() - t
        t = time.time()
        try:
            got = m.convolve(f, k, mode, 1.0)
            e = rel(got, ref.astype(np.float64)); mx = float(np.abs(got - ref).max())
            bad = np.argwhere(np.abs(got - ref) > 1e-3)
            print(shape, ks, mode, "rel_l2=%.3e max=%.3e gpu=%.3fs cpu=%.3fs" % (e, mx, time.time() - t, t1),
                  ("first bad zyx=%s n=%d" % (bad[0], len(bad))) if len(bad) else "", flush=True)
        except Exception as ex:
            print(shape, ks, mode, "FAILED", ex, flush=True)
