"""CPU tests (gloo, world_size 2) of the host logic of the sharded path: brick partition, ghost-layer
exchange, and that the two-exchanges-per-view schedule of ShardedDeconvolution reproduces the
single-brick result when the per-brick arithmetic is the oracle's.  No GPU, no library compute."""
import os
import socket

import numpy as np
import pytest

from spim_registration_b200.parallel import brick_grid, brick_of, rank_of


def test_brick_grid_and_bricks_partition_the_volume():
    for world in (1, 2, 4, 8):
        for dim in ((1024, 1024, 512), (2048, 2048, 1024), (130, 70, 50)):
            grid = brick_grid(world, dim)
            assert grid[0] * grid[1] * grid[2] == world
            cover = np.zeros(dim[::-1], np.int8) if dim[0] < 200 else None
            vol = 0
            for r in range(world):
                off, size, coords = brick_of(r, grid, dim)
                assert rank_of(coords, grid) == r
                assert all(s > 0 for s in size) and all(o % 4 == 0 for o in off[:1])
                vol += size[0] * size[1] * size[2]
                if cover is not None:
                    cover[off[2]:off[2] + size[2], off[1]:off[1] + size[1], off[0]:off[0] + size[0]] += 1
            assert vol == dim[0] * dim[1] * dim[2]
            if cover is not None:
                assert (cover == 1).all()
    assert brick_grid(8, (2048, 2048, 1024)) == (2, 2, 2)
    assert brick_grid(2, (1024, 1024, 512)) in ((1, 2, 1), (2, 1, 1))  # longest axis, y before x on ties


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, grid, q):
    import sys
    import torch
    import torch.distributed as dist
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import oracle_np as o
    from scipy import signal
    from spim_registration_b200.parallel import brick_of, exchange_ghosts
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dim = (24, 20, 16)  # x, y, z
        V, K, iters, lam = 2, 5, 2, 0.006
        rng = np.random.default_rng(11)
        imgs = [rng.random(dim[::-1], dtype=np.float32) + np.float32(0.05) for _ in range(V)]
        imgs[0][:, :, :3] = 0
        k1 = [o.norm_img(rng.random((K, K, K), dtype=np.float32), 8) for _ in range(V)]
        k2 = [o.compute_inverted_kernel(k) for k in k1]
        w = [np.float32(0.5)] * V
        ref, _, avg = o.run_deconvolution(imgs, w, k1, k2, iters, lam)

        off, size, coords = brick_of(rank, grid, dim)
        h = K // 2
        ghost = tuple(h if grid[d] > 1 else 0 for d in range(3))
        glo = [min(ghost[d], off[d]) for d in range(3)]
        ghi = [min(ghost[d], dim[d] - off[d] - size[d]) for d in range(3)]
        adim = [size[d] + glo[d] + ghi[d] for d in range(3)]
        sl = tuple(slice(off[d], off[d] + size[d]) for d in (2, 1, 0))
        inner = tuple(slice(glo[d], glo[d] + size[d]) for d in (2, 1, 0))

        def ghosted(fill):
            return torch.full(tuple(adim[::-1]), float(fill), dtype=torch.float32)

        def exchange(t):
            exchange_ghosts(t, (glo[2], glo[1], glo[0]), size[::-1], ghost, coords, grid, dist)

        def conv_brick(t, kernel, mode, value=0.0):
            # ghost layers supply the neighbours' voxels, the reference's rule applies at the volume border
            a = t.numpy().astype(np.float64)
            lo = [0 if glo[d] else kernel.shape[2 - d] - 1 - kernel.shape[2 - d] // 2 for d in range(3)]
            hi = [0 if ghi[d] else kernel.shape[2 - d] // 2 for d in range(3)]
            ext = o.extend(a, lo[::-1], hi[::-1], mode, value)
            full = signal.fftconvolve(ext, kernel.astype(np.float64), mode="valid")
            # 'valid' output starts at array index (K-1-K//2 - lo) ... crop to the brick
            s = [glo[d] - ((kernel.shape[2 - d] - 1 - kernel.shape[2 - d] // 2) - lo[d]) for d in range(3)]
            return full[s[2]:s[2] + size[2], s[1]:s[1] + size[1], s[0]:s[0] + size[0]].astype(np.float32)

        psi = ghosted(avg)
        quot = ghosted(1.0)
        for _ in range(iters):
            for v in range(V):
                t1 = conv_brick(psi, k1[v], "mirror")
                quot[inner] = torch.from_numpy(o.compute_quotient(t1, imgs[v][sl]))
                exchange(quot)
                t2 = conv_brick(quot, k2[v], "const", 1.0)
                new, _, _ = o.compute_final_values(psi[inner].numpy(), t2, w[v], lam)
                psi[inner] = torch.from_numpy(new)
                exchange(psi)
        out = psi[inner].numpy().copy()
        cnt = sum((im[sl] > 0).astype(np.int32) for im in imgs)
        out[cnt == 0] = 0
        err = float(np.abs(out - ref[sl]).max() / np.abs(ref).max())
        q.put((rank, err))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("grid", [(2, 1, 1), (1, 1, 2)])
def test_sharded_schedule_equals_single_brick_gloo(grid):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, grid, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err in res:
        assert err <= 2e-6, (rank, err)


def test_brick_grid_by_window_cost():
    """With the kernel extent given the grid minimises the planner's FFT-window volume of the largest brick
    (parallel.axis_window_cost mirrors csrc/engine.cu plan_axis); without it the round-1 rule (longest axis first)."""
    from spim_registration_b200.parallel import XY_WINDOWS, Z_WINDOWS, axis_window_cost, brick_grid, brick_of
    assert axis_window_cost(1024, 31, 4, XY_WINDOWS) == 11 * (128 + 6)                # eleven 128-point windows of 96 outputs each
    assert axis_window_cost(512, 31, 1, Z_WINDOWS) == (512 + 6) + (64 + 6)
    assert brick_grid(8, (1024, 1024, 512)) == (2, 2, 2)
    assert brick_grid(8, (1024, 1024, 512), (31, 31, 31)) == (2, 4, 1)                # 1.93 x window inflation instead of 2.36 x
    assert brick_grid(8, (2048, 2048, 1024), (31, 31, 31)) == (1, 4, 2)
    assert brick_grid(1, (1024, 1024, 512), (31, 31, 31)) == (1, 1, 1)
    for world in (2, 4, 8):
        g = brick_grid(world, (192, 160, 96), (15, 15, 15))
        assert g[0] * g[1] * g[2] == world
        seen = set()
        for r in range(world):                                                       # bricks tile the volume exactly
            off, size, _ = brick_of(r, g, (192, 160, 96))
            seen.add((off, size))
            assert all(s >= 7 for s in size)
        assert sum(s[0] * s[1] * s[2] for _, s in seen) == 192 * 160 * 96
    # a volume too thin to be cut along z for this kernel: no grid cuts it there
    assert brick_grid(8, (1024, 1024, 20), (31, 31, 31))[2] == 1
