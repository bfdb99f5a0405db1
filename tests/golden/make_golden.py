"""Generates tests/golden/loop_3view_32.npz with the numpy oracle (oracle/oracle_np.py).

The reference has no golden vectors for this path, so the fixture freezes
the oracle's own output: 3 views, 24x28x32 voxels, extracted-style PSFs 7^3, efficient Bayesian,
5 iterations, lambda 0.006, blending weights.  Run from the repo root: python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_np as o  # noqa: E402
from spim_registration_b200 import synthetic  # noqa: E402 (numpy-only module)


def main():
    shape, V, iters, lam, psftype = (24, 28, 32), 3, 5, 0.006, "EFFICIENT_BAYESIAN"
    ds = synthetic.make_dataset(shape, V, (7, 7, 7), extracted=True, seed=synthetic.SEED0 + 77,
                                convolve=lambda f, k: o.convolve_extended(f, k, "mirror"))
    weights = synthetic.blending_weights(shape, V, border=(3, 3, 1), blend_range=6.0)
    k1, k2 = o.init_kernels(ds["psfs"], psftype, 8)
    psi, stats, avg = o.run_deconvolution(ds["imgs"], weights, k1, k2, iters, lam)
    out = dict(psi=psi, avg=np.float64(avg), num_views=np.int32(V), iterations=np.int32(iters), lam=np.float64(lam),
               psftype=np.array(psftype), sum_change=np.array([[s for s, _ in r] for r in stats]),
               max_change=np.array([[m for _, m in r] for r in stats]))
    for v in range(V):
        out["img%d" % v] = ds["imgs"][v]
        out["weight%d" % v] = weights[v]
        out["psf%d" % v] = ds["psfs"][v]
        out["k1_%d" % v] = k1[v]
        out["k2_%d" % v] = k2[v]
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "loop_3view_32.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; avg", avg)


if __name__ == "__main__":
    main()
