"""Host-side cost of setting a brick up from host buffers (what the end-to-end leg of an 8-GPU run pays per rank), on one GPU.
usage: python scripts/prof_setup.py [world=8] [rank=3] [config=3]"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import spim_registration_b200 as m  # noqa: E402
from spim_registration_b200 import _lib as L  # noqa: E402
from spim_registration_b200.parallel import ShardedDeconvolution, brick_grid, brick_of  # noqa: E402


class FakeDist:
    def __init__(self, world, rank):
        self.w, self.r = world, rank

    def get_world_size(self):
        return self.w

    def get_rank(self):
        return self.r

    def all_reduce(self, t, op=None):
        return None


def main():
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    rank = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    cfg = bench.CONFIGS[int(sys.argv[3]) if len(sys.argv) > 3 else 3]
    torch.cuda.set_device(0)
    lib = m.load()
    Z, Y, X = cfg["shape"]
    dim = (X, Y, Z)
    psfs, k1, k2 = bench.prepare_kernels(cfg)
    grid = brick_grid(world, dim, [cfg["k"]] * 3)
    off, size, _ = brick_of(rank, grid, dim)
    V = cfg["views"]
    h_imgs = [torch.rand(size[::-1], dtype=torch.float32).pin_memory() for _ in range(V)]
    np_imgs = [h.numpy() for h in h_imgs]
    w = [torch.full(size[::-1], 1.0 / V, device="cuda") for _ in range(V)]
    torch.cuda.synchronize()
    for rep in range(3):
        t0 = time.perf_counter()
        dec = ShardedDeconvolution(dim, np_imgs, w, k1, k2, bench.LAMBDA, FakeDist(world, rank), 0, grid=grid, async_upload=True)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        dec.init_psi_average()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        dec.close()
        t4 = time.perf_counter()
        print("rep %d: create+add_views %.1f ms (uploads drained after another %.1f ms), psi init %.1f ms, close %.1f ms" %
              (rep, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3)))
    # the pieces of one creation
    os.environ["MVD_TRACE_SETUP"] = "1"
    ctx = C.c_void_p()
    gh = [15 if grid[d] > 1 else 0 for d in range(3)]
    t0 = time.perf_counter()
    L.check(lib.mvdCreate(C.byref(ctx), 0, L.int3(dim), L.int3(off), L.int3(size), L.int3(gh), bench.LAMBDA, 1e-4), "mvdCreate")
    t1 = time.perf_counter()
    print("mvdCreate %.1f ms" % (1e3 * (t1 - t0)))
    for v in range(V):
        a = np.ascontiguousarray(k1[v], dtype=np.float32)
        b = np.ascontiguousarray(k2[v], dtype=np.float32)
        t0 = time.perf_counter()
        L.check(lib.mvdAddView(ctx, np_imgs[v].ctypes.data_as(C.c_void_p), C.c_void_p(w[v].data_ptr()), 1.0, a.ctypes.data_as(L.c_float_p), L.int3(a.shape[::-1]),
                               b.ctypes.data_as(L.c_float_p), L.int3(b.shape[::-1]), 4, None), "mvdAddView")
        t1 = time.perf_counter()
        print("mvdAddView %d: %.1f ms" % (v, 1e3 * (t1 - t0)))
    torch.cuda.synchronize()
    lib.mvdDestroy(ctx)


if __name__ == "__main__":
    main()
