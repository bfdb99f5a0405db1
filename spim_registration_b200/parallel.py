"""Spatial sharding of the deconvolution over the GPUs of one box.

One process per GPU (torch.distributed, NCCL over NVLink; gloo for the CPU tests).  The volume is
cut into a grid of bricks; every rank keeps psi and the quotient for its brick plus a ghost layer
of the kernel half-width.  Per view and iteration there are exactly two exchanges -- the quotient
ghosts before conv2 and the psi ghosts before the next conv1 -- done axis by axis (x, y, z) with
full-extent slabs so that edges and corners arrive through the later axes.  The reference has no
counterpart (its multi-GPU mode is one Java thread per device pulling blocks through host RAM,
MVDeconFFTThreads.java:96-134).
"""
import ctypes as C

import numpy as np


XY_WINDOWS = (32, 64, 128)
Z_WINDOWS = (16, 32, 64, 128, 256, 512)


def axis_window_cost(n, k, align, sizes):
    """Total length of the overlap-save windows the library's planner (csrc/engine.cu plan_axis: dynamic programme over the
    first uncovered voxel, cost = window length + 6 per window) uses to cover n outputs with a kernel of extent k."""
    hl, hr = k - 1 - k // 2, k // 2
    lo = (hl + align - 1) // align * align
    eff = [(s - lo - hr) // align * align for s in sizes]
    if n <= 0 or not any(e > 0 for e in eff):
        return float("inf")
    cost = [0] * (n + 1)
    for pos in range(n - 1, -1, -1):
        cost[pos] = min(s + 6 + cost[min(n, pos + e)] for s, e in zip(sizes, eff) if e > 0)
    return cost[0]


def _axis_bounds(g, n):
    return [0 if i == 0 else (n if i == g else (n * i) // g // 4 * 4) for i in range(g + 1)]


def brick_grid(world, dim_xyz, kernel_xyz=None):
    """Grid (gx, gy, gz) with gx*gy*gz = world.  With the kernel extent given: the grid whose largest brick needs the
    least window volume (bricks whose extents sit badly on the FFT windows cost up to 20 % more work: for 1024 x 1024 x
    512, K = 31 on 8 GPUs 2 x 2 x 2 inflates 2.36 x, 2 x 4 x 1 only 1.93 x, like the undivided volume); ties prefer
    fewer cuts along x (long rows), then along y.  Without it: split the longest axes first, z before y before x on ties."""
    n = world
    assert n >= 1 and (n & (n - 1)) == 0, "world size must be a power of two"
    if kernel_xyz is not None:
        best, best_key = None, None
        gx = 1
        while gx <= world:
            gy = 1
            while gx * gy <= world:
                gz = world // (gx * gy)
                grid = (gx, gy, gz)
                cost = 1.0
                ok = True
                for d in range(3):
                    b = _axis_bounds(grid[d], dim_xyz[d])
                    sizes = [b[i + 1] - b[i] for i in range(grid[d])]
                    ghost = max(kernel_xyz[d] // 2, kernel_xyz[d] - 1 - kernel_xyz[d] // 2) if grid[d] > 1 else 0
                    if min(sizes) < max(1, ghost):
                        ok = False
                        break
                    cost *= max(axis_window_cost(sz, kernel_xyz[d], 4 if d < 2 else 1, XY_WINDOWS if d < 2 else Z_WINDOWS) for sz in sizes)
                if ok and cost != float("inf"):
                    key = (cost, gx, gy)
                    if best_key is None or key < best_key:
                        best, best_key = grid, key
                gy *= 2
            gx *= 2
        if best is not None:
            return best
    grid = [1, 1, 1]
    while n > 1:
        per = [dim_xyz[d] / grid[d] for d in range(3)]
        d = max((2, 1, 0), key=lambda a: (per[a], a))
        grid[d] *= 2
        n //= 2
    return tuple(grid)


def brick_of(rank, grid, dim_xyz):
    """(offset_xyz, size_xyz, coords) of `rank`; x fastest in the rank order."""
    cx, cy, cz = rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1])
    # brick boundaries b_i = floor4(n*i/g) (b_0 = 0, b_g = n) keep x bricks float4-aligned
    off, size = [], []
    for c, g, n in zip((cx, cy, cz), grid, dim_xyz):
        def bound(i):
            if i == 0:
                return 0
            if i == g:
                return n
            return (n * i) // g // 4 * 4
        off.append(bound(c))
        size.append(bound(c + 1) - bound(c))
    return tuple(off), tuple(size), (cx, cy, cz)


def rank_of(coords, grid):
    return coords[0] + grid[0] * (coords[1] + grid[1] * coords[2])


def exchange_ghosts(t, off_zyx, brick_zyx, ghost_xyz, coords, grid, dist, group=None):
    """Fill the ghost layers of the ghosted tensor `t` ([z,y,x], brick interior starting at
    off_zyx) from the neighbouring ranks.  Works for CUDA tensors (NCCL) and CPU tensors (gloo)."""
    import torch
    for axis_xyz in (0, 1, 2):
        g = ghost_xyz[axis_xyz]
        if g == 0 or grid[axis_xyz] == 1:
            continue
        dim = 2 - axis_xyz  # tensor dim
        o, b = off_zyx[dim], brick_zyx[dim]
        c = coords[axis_xyz]
        ops, recvs = [], []
        for side in (-1, 1):
            nc = c + side
            if nc < 0 or nc >= grid[axis_xyz]:
                continue
            nco = list(coords)
            nco[axis_xyz] = nc
            peer = rank_of(nco, grid)
            if side < 0:
                send = t.narrow(dim, o, g).contiguous()        # my low interior -> peer's high ghost
                ghost = t.narrow(dim, o - g, g)                 # my low ghost <- peer's high interior
            else:
                send = t.narrow(dim, o + b - g, g).contiguous()
                ghost = t.narrow(dim, o + b, g)
            rbuf = torch.empty(ghost.shape, dtype=t.dtype, device=t.device)
            ops.append(dist.P2POp(dist.isend, send, peer, group))
            ops.append(dist.P2POp(dist.irecv, rbuf, peer, group))
            recvs.append((ghost, rbuf))
        if ops:
            for r in dist.batch_isend_irecv(ops):
                r.wait()
            for ghost, rbuf in recvs:
                ghost.copy_(rbuf)


class _CudaArray:
    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = dict(shape=tuple(shape), typestr="<f4", data=(int(ptr), False), version=2)


class ShardedDeconvolution:
    """MVDeconvolution (MVDeconvolution.java:73-190) on one brick per rank through the step API of
    libmvdecon.so.  imgs/weights are this rank's brick ([z,y,x] torch CUDA tensors or numpy arrays)."""

    def __init__(self, dim_xyz, imgs, weights, kernel1, kernel2, lam, dist=None, device=0, min_value=1e-4, grid=None, async_upload=False):
        """async_upload: host images (pinned, kept alive by the caller until the first run_iterations / stats call has
        returned) are uploaded on the library's copy stream while the caller goes on: MVD_MEM_HOST_ASYNC."""
        import torch
        from . import _lib as L
        self.L, self.lib, self.torch, self.dist = L, L.load(), torch, dist
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0
        self.dim = tuple(dim_xyz)
        self.grid = grid or brick_grid(self.world, self.dim)
        self.off, self.size, self.coords = brick_of(self.rank, self.grid, self.dim)
        self.device = device
        self.V = len(imgs)
        gh = [0, 0, 0]
        for k in list(kernel1) + list(kernel2):
            ks = k.shape[::-1]
            for d in range(3):
                if self.grid[d] > 1:
                    gh[d] = max(gh[d], ks[d] // 2, ks[d] - 1 - ks[d] // 2)
        self.ghost = tuple(gh)
        self.ctx = C.c_void_p()
        L.check(self.lib.mvdCreate(C.byref(self.ctx), device, L.int3(self.dim), L.int3(self.off), L.int3(self.size), L.int3(self.ghost),
                                   float(lam), float(min_value)), "mvdCreate")
        L.check(self.lib.mvdSetStream(self.ctx, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "mvdSetStream")
        for v in range(self.V):
            img, w = imgs[v], weights[v]
            is_dev = hasattr(img, "data_ptr")
            ip = C.c_void_p(img.data_ptr()) if is_dev else img.ctypes.data_as(C.c_void_p)
            w_dev = hasattr(w, "data_ptr")
            if w is None or np.isscalar(w):
                wp, wc = None, (1.0 if w is None else float(w))
            else:
                assert w_dev or not is_dev, "device images with host weights are not supported"
                wp, wc = (C.c_void_p(w.data_ptr()) if w_dev else w.ctypes.data_as(C.c_void_p)), 1.0
            k1 = np.ascontiguousarray(kernel1[v], dtype=np.float32)
            k2 = np.ascontiguousarray(kernel2[v], dtype=np.float32)
            L.check(self.lib.mvdAddView(self.ctx, ip, wp, wc, k1.ctypes.data_as(L.c_float_p), L.int3(k1.shape[::-1]),
                                        k2.ctypes.data_as(L.c_float_p), L.int3(k2.shape[::-1]),
                                        1 if is_dev else ((4 if w_dev else 3) if async_upload else (2 if w_dev else 0)), None), "mvdAddView")
        self._arrays = {}
        for which in (0, 1):
            ptr, dim, off = C.c_void_p(), (C.c_int * 3)(), (C.c_int * 3)()
            L.check(self.lib.mvdArrayInfo(self.ctx, which, C.byref(ptr), dim, off), "mvdArrayInfo")
            t = torch.as_tensor(_CudaArray(ptr.value, (dim[2], dim[1], dim[0])), device=torch.device("cuda", device))
            self._arrays[which] = (t, (off[2], off[1], off[0]))
        self.slot = 0
        self.fused = False

    # -- fused ghost exchange ---------------------------------------------------------------------
    def neighbours(self):
        """[(rank, offset_xyz, size_xyz)] of the bricks whose ghost layers overlap this brick."""
        out = []
        for dz in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    delta = (dx, dy, dz)
                    if delta == (0, 0, 0) or any(delta[d] and not self.ghost[d] for d in range(3)):
                        continue
                    nc = [self.coords[d] + delta[d] for d in range(3)]
                    if any(nc[d] < 0 or nc[d] >= self.grid[d] for d in range(3)):
                        continue
                    r = rank_of(nc, self.grid)
                    off, size, _ = brick_of(r, self.grid, self.dim)
                    out.append((r, off, size))
        return out

    def _slot(self, coords):
        """Slot of the brick at `coords` in THIS brick's flag array: the 3 x 3 x 3 code of its offset."""
        d = [coords[i] - self.coords[i] for i in range(3)]
        return (d[0] + 1) + 3 * (d[1] + 1) + 9 * (d[2] + 1)

    def attach_peers(self):
        """Switch to the fused exchange (mvdAttachPeer, include/mvdecon.h): the step kernels store
        ghost voxels straight into the neighbours' arrays over NVLink; between the steps only a
        barrier remains.  One process per GPU: the arrays are shared through CUDA IPC handles.
        The barrier is neighbour-only, in peer memory (mvdPeerBarrier: every rank stores the step number into its
        neighbours' flag arrays and waits for theirs); MVD_BARRIER=nccl keeps round 1's 4-byte all-reduce."""
        if self.world == 1:
            return
        import os
        torch, L = self.torch, self.L
        for d in range(3):
            assert self.grid[d] == 1 or self.size[d] >= self.ghost[d], "bricks thinner than the ghost layer are not supported"
        hp, hq, hf = (C.c_ubyte * 64)(), (C.c_ubyte * 64)(), (C.c_ubyte * 64)()
        L.check(self.lib.mvdExportArrays(self.ctx, hp, hq), "mvdExportArrays")
        L.check(self.lib.mvdExportFlags(self.ctx, hf, None), "mvdExportFlags")
        dev = self._arrays[0][0].device
        mine = torch.tensor(list(bytes(hp) + bytes(hq) + bytes(hf)), dtype=torch.uint8, device=dev)
        every = [torch.empty_like(mine) for _ in range(self.world)]
        self.dist.all_gather(every, mine)
        self.flag_barrier = os.environ.get("MVD_BARRIER", "flags") == "flags"
        for r, off, size in self.neighbours():
            hb = bytes(every[r].cpu().numpy().tobytes())
            a, b = C.create_string_buffer(hb[:64], 64), C.create_string_buffer(hb[64:128], 64)
            L.check(self.lib.mvdAttachPeer(self.ctx, L.int3(off), L.int3(size), a, b, 1, -1), "mvdAttachPeer")
            if self.flag_barrier:
                nc = (r % self.grid[0], (r // self.grid[0]) % self.grid[1], r // (self.grid[0] * self.grid[1]))
                mine_in_theirs = 26 - self._slot(nc)   # the code of the opposite offset
                L.check(self.lib.mvdAttachPeerFlags(self.ctx, C.create_string_buffer(hb[128:192], 64), 1, mine_in_theirs, self._slot(nc)),
                        "mvdAttachPeerFlags")
        self._flag = torch.zeros(1, dtype=torch.float32, device=dev)
        self._barrier_count = 0
        self.fused = True
        self.dist.barrier()

    # -- pieces of the MVDeconvolution constructor ------------------------------------------
    def init_psi_average(self):
        s, n = C.c_double(), C.c_longlong()
        self.L.check(self.lib.mvdFirstIterationSums(self.ctx, C.byref(s), C.byref(n)), "mvdFirstIterationSums")
        tot = self.torch.tensor([s.value, float(n.value)], dtype=self.torch.float64, device=self._arrays[0][0].device)
        if self.world > 1:
            self.dist.all_reduce(tot)
        avg = float(tot[0] / tot[1])
        if avg != avg:
            avg = 0.5
        self.L.check(self.lib.mvdFillPsi(self.ctx, avg), "mvdFillPsi")
        return avg

    def set_psi(self, brick):
        """mvdSetPsi for this rank's brick ([z,y,x] numpy or CUDA tensor) followed by a real ghost exchange: in
        fused mode `_exchange` is only a barrier (the step kernels fill the ghosts), so a psi that did not come
        from a step has to be sent explicitly."""
        is_dev = hasattr(brick, "data_ptr")
        ptr = C.c_void_p(brick.data_ptr()) if is_dev else np.ascontiguousarray(brick, dtype=np.float32).ctypes.data_as(C.c_void_p)
        self.L.check(self.lib.mvdSetPsi(self.ctx, ptr, 1 if is_dev else 0), "mvdSetPsi")
        if self.world > 1:
            t, off = self._arrays[0]
            exchange_ghosts(t, off, self.size[::-1], self.ghost, self.coords, self.grid, self.dist)
            if self.fused:
                self.dist.all_reduce(self._flag)   # rare (once per mvdSetPsi): a collective is fine here

    def _exchange(self, which):
        if self.world == 1:
            return
        if self.fused:
            # the ghost layers were written by the neighbours' step kernels; every NEIGHBOUR must have
            # finished the step before the next one reads them (stream-ordered, no host sync)
            if getattr(self, "flag_barrier", False):
                self._barrier_count += 1
                self.L.check(self.lib.mvdPeerBarrier(self.ctx, self._barrier_count), "mvdPeerBarrier")
            else:
                self.dist.all_reduce(self._flag)
            return
        t, off = self._arrays[which]
        exchange_ghosts(t, off, self.size[::-1], self.ghost, self.coords, self.grid, self.dist)

    def run_iterations(self, n):
        first = self.slot
        if self.world == 1:
            for _ in range(n):
                for v in range(self.V):
                    self.L.check(self.lib.mvdStepQuotient(self.ctx, v), "mvdStepQuotient")
                    self.L.check(self.lib.mvdStepUpdate(self.ctx, v, self.slot), "mvdStepUpdate")
                    self.slot += 1
            return first, self.slot
        for _ in range(n):
            for v in range(self.V):
                self.L.check(self.lib.mvdStepQuotient(self.ctx, v), "mvdStepQuotient")
                self._exchange(1)
                self.L.check(self.lib.mvdStepUpdate(self.ctx, v, self.slot), "mvdStepUpdate")
                self._exchange(0)
                self.slot += 1
        return first, self.slot

    def stats(self, first, last):
        n = last - first
        s, m = (C.c_double * n)(), (C.c_double * n)()
        self.L.check(self.lib.mvdGetStats(self.ctx, first, n, s, m), "mvdGetStats")
        ts = self.torch.tensor(list(s), dtype=self.torch.float64, device=self._arrays[0][0].device)
        tm = self.torch.tensor(list(m), dtype=self.torch.float64, device=self._arrays[0][0].device)
        if self.world > 1:
            self.dist.all_reduce(ts)
            self.dist.all_reduce(tm, op=self.dist.ReduceOp.MAX)
        return ts.cpu().numpy().reshape(-1, self.V), tm.cpu().numpy().reshape(-1, self.V)

    def mask_never_updated(self):
        self.L.check(self.lib.mvdMaskNeverUpdated(self.ctx), "mvdMaskNeverUpdated")

    def get_psi_brick(self, out=None):
        """Brick of psi as numpy [z,y,x] (device -> host)."""
        if out is None:
            out = np.empty(self.size[::-1], dtype=np.float32)
        self.L.check(self.lib.mvdGetPsi(self.ctx, out.ctypes.data_as(C.c_void_p), 0), "mvdGetPsi")
        return out

    def kernel_launches(self):
        return self.lib.mvdKernelLaunchCount()

    def close(self):
        if self.ctx:
            if self.fused and self.world > 1 and hasattr(self.dist, "barrier"):
                self.torch.cuda.synchronize()
                self.dist.barrier()  # no neighbour may still be writing into arrays that are about to be freed
            self.lib.mvdDestroy(self.ctx)
            self.ctx = None


class LocalBrickGroup:
    """All bricks of a grid on ONE device, stepped in lock-step with in-process ghost copies -- the
    same library calls and the same exchange schedule as the multi-process path, without
    torch.distributed.  Used by the single-GPU parity test of the brick/ghost code paths."""

    def __init__(self, dim_xyz, grid, imgs, weights, kernel1, kernel2, lam, device=0, use_flags=False):
        import torch
        self.torch = torch
        self.use_flags = use_flags
        self.grid, self.dim = tuple(grid), tuple(dim_xyz)
        self.world = grid[0] * grid[1] * grid[2]
        self.bricks = []
        for r in range(self.world):
            off, size, _ = brick_of(r, self.grid, self.dim)
            sl = tuple(slice(off[d], off[d] + size[d]) for d in (2, 1, 0))
            bi = [np.ascontiguousarray(i[sl]) for i in imgs]
            bw = [w if (w is None or np.isscalar(w)) else np.ascontiguousarray(w[sl]) for w in weights]
            d = ShardedDeconvolution.__new__(ShardedDeconvolution)
            _init_brick(d, self.dim, bi, bw, kernel1, kernel2, lam, device, self.grid, r, self.world)
            self.bricks.append(d)
        self.fused = False

    def attach_peers(self):
        """Fused exchange between bricks of one device: neighbours are plain device pointers."""
        for b in self.bricks:
            for r, off, size in b.neighbours():
                peer = self.bricks[r]
                pp, pq = peer._arrays[0][0].data_ptr(), peer._arrays[1][0].data_ptr()
                b.L.check(b.lib.mvdAttachPeer(b.ctx, b.L.int3(off), b.L.int3(size), C.c_void_p(pp), C.c_void_p(pq), 0, -1), "mvdAttachPeer")
                if self.use_flags:
                    fp = C.c_void_p()
                    peer.L.check(peer.lib.mvdExportFlags(peer.ctx, None, C.byref(fp)), "mvdExportFlags")
                    b.L.check(b.lib.mvdAttachPeerFlags(b.ctx, fp, 0, 26 - b._slot(peer.coords), b._slot(peer.coords)), "mvdAttachPeerFlags")
        self.fused = True
        self._barrier_count = 0

    def _exchange(self, which):
        if self.fused:
            # every brick runs on the same stream: the steps are already ordered.  With use_flags the neighbour-only flag
            # barrier of the multi-process path is exercised as well -- ALL bricks signal before ANY brick waits (kernels of
            # several bricks on one GPU that wait for one another may never run at the same time)
            if self.use_flags:
                self._barrier_count += 1
                for b in self.bricks:
                    if b.neighbours():
                        b.L.check(b.lib.mvdPeerSignal(b.ctx, self._barrier_count), "mvdPeerSignal")
                for b in self.bricks:
                    if b.neighbours():
                        b.L.check(b.lib.mvdPeerWait(b.ctx, self._barrier_count), "mvdPeerWait")
                for b in self.bricks:
                    b.L.check(b.lib.mvdSynchronize(b.ctx), "mvdSynchronize")   # raises if a bounded spin expired
            return
        # axis by axis, full-extent slabs (edges and corners travel through the later axes)
        for axis in (0, 1, 2):
            dimt = 2 - axis
            for b in self.bricks:
                g = b.ghost[axis]
                if g == 0 or self.grid[axis] == 1:
                    continue
                t, off = b._arrays[which]
                o, n = off[dimt], b.size[axis]
                for side in (-1, 1):
                    nc = list(b.coords)
                    nc[axis] += side
                    if nc[axis] < 0 or nc[axis] >= self.grid[axis]:
                        continue
                    peer = self.bricks[rank_of(nc, self.grid)]
                    pt, poff = peer._arrays[which]
                    po, pn = poff[dimt], peer.size[axis]
                    if side < 0:
                        t.narrow(dimt, o - g, g).copy_(pt.narrow(dimt, po + pn - g, g))
                    else:
                        t.narrow(dimt, o + n, g).copy_(pt.narrow(dimt, po, g))

    def init_psi_average(self):
        import ctypes as C
        tot, cnt = 0.0, 0
        for b in self.bricks:
            s, n = C.c_double(), C.c_longlong()
            b.L.check(b.lib.mvdFirstIterationSums(b.ctx, C.byref(s), C.byref(n)), "mvdFirstIterationSums")
            tot += s.value
            cnt += n.value
        avg = tot / cnt if cnt else 0.5
        for b in self.bricks:
            b.L.check(b.lib.mvdFillPsi(b.ctx, avg), "mvdFillPsi")
        return avg

    def run_iterations(self, n):
        V = self.bricks[0].V
        for _ in range(n):
            for v in range(V):
                for b in self.bricks:
                    b.L.check(b.lib.mvdStepQuotient(b.ctx, v), "mvdStepQuotient")
                self.torch.cuda.synchronize()
                self._exchange(1)
                for b in self.bricks:
                    b.L.check(b.lib.mvdStepUpdate(b.ctx, v, b.slot), "mvdStepUpdate")
                    b.slot += 1
                self.torch.cuda.synchronize()
                self._exchange(0)

    def get_psi(self):
        out = np.empty(self.dim[::-1], dtype=np.float32)
        for b in self.bricks:
            b.mask_never_updated()
            sl = tuple(slice(b.off[d], b.off[d] + b.size[d]) for d in (2, 1, 0))
            out[sl] = b.get_psi_brick()
        return out

    def close(self):
        for b in self.bricks:
            b.close()


def _init_brick(d, dim_xyz, imgs, weights, kernel1, kernel2, lam, device, grid, rank, world):
    """ShardedDeconvolution.__init__ for an explicit (rank, world) without a process group."""
    class _FakeDist:
        def get_world_size(self):
            return world

        def get_rank(self):
            return rank
    ShardedDeconvolution.__init__(d, dim_xyz, imgs, weights, kernel1, kernel2, lam, _FakeDist(), device, grid=grid)
