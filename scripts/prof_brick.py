"""One brick of a multi-GPU run on a single device, without neighbours: what rank `r` of an N-GPU
bench executes per step (kernel times of the brick/ghost code paths), minus the exchange.

usage: python scripts/prof_brick.py [world=8] [rank=0] [steps=3] [config=3]
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402
import spim_registration_b200 as m  # noqa: E402
from spim_registration_b200 import _lib as L  # noqa: E402
from spim_registration_b200.parallel import ShardedDeconvolution, _init_brick, brick_grid, brick_of  # noqa: E402


def main():
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    rank = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    cfg = bench.CONFIGS[int(sys.argv[4]) if len(sys.argv) > 4 else 3]
    torch.cuda.set_device(0)
    lib = m.load()
    Z, Y, X = cfg["shape"]
    dim = (X, Y, Z)
    psfs, k1, k2 = bench.prepare_kernels(cfg)
    g = os.environ.get("MVD_GRID", "cost")   # "cost" (default rule), "longest" (round 1's rule) or an explicit "gx,gy,gz"
    grid = brick_grid(world, dim, [cfg["k"]] * 3) if g == "cost" else (brick_grid(world, dim) if g == "longest" else tuple(int(x) for x in g.split(",")))
    off, size, _ = brick_of(rank, grid, dim)
    imgs, weights = bench.gen_dataset_device(cfg, lib, psfs, (off[::-1], size[::-1]))
    dec = ShardedDeconvolution.__new__(ShardedDeconvolution)
    _init_brick(dec, dim, imgs, weights, k1, k2, bench.LAMBDA, 0, grid, rank, world)
    dec.lib.mvdFillPsi(dec.ctx, 1.0)
    if os.environ.get("MVD_FAKE_PEERS", "0") == "1":
        # the push code path without a second GPU: every neighbour's arrays are dummies on this device
        dummies = []
        for r, noff, nsize in dec.neighbours():
            n = (nsize[0] + 32) * (nsize[1] + 30) * (nsize[2] + 30)
            a, b = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
            dummies.append((a, b))
            L.check(lib.mvdAttachPeer(dec.ctx, L.int3(noff), L.int3(nsize), C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), 0, -1), "mvdAttachPeer")
        print("attached %d fake peers" % len(dummies))
    V = dec.V

    def run(n):
        for _ in range(n):
            for v in range(V):
                L.check(lib.mvdStepQuotient(dec.ctx, v), "q")
                L.check(lib.mvdStepUpdate(dec.ctx, v, dec.slot), "u")
                dec.slot += 1

    run(2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    L.check(lib.mvdProfileStart(dec.ctx), "start")
    run(1)
    torch.cuda.synchronize()
    pms, pl, pb = (C.c_double * 3)(), (C.c_longlong * 3)(), (C.c_double * 3)()
    L.check(lib.mvdProfileStop(dec.ctx, pms, pl, pb), "stop")
    names = ("fwd_xy", "z_pass", "inv_xy")
    out = {"grid": grid, "rank": rank, "brick_off": off, "brick_size": size, "ghost": dec.ghost,
           "array_dim_xyz": list(dec._arrays[0][0].shape[::-1]), "array_off_xyz": list(dec._arrays[0][1][::-1]),
           "ms_per_step": ms, "brick_gvoxel_it_s": size[0] * size[1] * size[2] / ms / 1e6,
           "kernels_serial_ms_per_step": {names[i]: pms[i] for i in range(3)},
           "kernels_gbs": {names[i]: pb[i] / 1e6 / pms[i] for i in range(3)}}
    print(json.dumps(out))
    dec.close()


if __name__ == "__main__":
    main()
