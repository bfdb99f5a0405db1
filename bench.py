#!/usr/bin/env python
"""bench.py -- Gvoxel-iterations/s of the multi-view deconvolution loop on N B200s.

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--config 3]

A step is one iteration of the loop (all views: conv1 -> quotient -> conv2 -> update) on the
synthetic workload of BASELINE.json configs[2] (7 views, 1024x1024x512, PSF 31^3, efficient
Bayesian), the configuration the north_star target is quoted on; it fits one GPU.  One JSON line
is printed by rank 0 (see the task contract).  `--impl reference` times the CPU port of the
reference's CPU path (oracle/cpu_baseline.py) on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

CONFIGS = {
    # shape is (z, y, x)
    1: dict(name="cfg1: 3-view 128^3 blob volume, Gaussian PSF 15^3, efficient Bayesian", views=3, shape=(128, 128, 128), k=15,
            psftype="EFFICIENT_BAYESIAN", kind="blob", extracted=False, dense_weights=False),
    2: dict(name="cfg2: 6-view 512x512x256 bead phantom, extracted PSF 31^3, OptimizationII", views=6, shape=(256, 512, 512), k=31,
            psftype="OPTIMIZATION_II", kind="bead", extracted=True, dense_weights=False),
    3: dict(name="cfg3: 7-view 1024x1024x512 volume, PSF 31^3, efficient Bayesian, per-view blending weights resident in HBM", views=7,
            shape=(512, 1024, 1024), k=31, psftype="EFFICIENT_BAYESIAN", kind="blob", extracted=False, dense_weights=True),
    4: dict(name="cfg4: 6-view 1024x1024x512 independent RL with per-view weight images", views=6, shape=(512, 1024, 1024), k=31,
            psftype="INDEPENDENT", kind="blob", extracted=False, dense_weights=True),
    5: dict(name="cfg5: 8-view 2048x2048x1024 volume sharded by bricks across 8 GPUs (2x2x2), PSF 31^3, efficient Bayesian, per-view blending weights",
            views=8, shape=(1024, 2048, 2048), k=31, psftype="EFFICIENT_BAYESIAN", kind="blob", extracted=False, dense_weights=True, min_gpus=8),
    # debugging size
    0: dict(name="cfg0: 3-view 96x160x192 debug volume, PSF 15^3", views=3, shape=(96, 160, 192), k=15,
            psftype="EFFICIENT_BAYESIAN", kind="blob", extracted=False, dense_weights=True),
}
LAMBDA = 0.006
ALG_BYTES_PER_VOXEL_VIEW = {True: 36.0, False: 32.0}  # SURVEY.md 8d / BASELINE.md section 3


_REAL_STDOUT = None


def quiet_stdout():
    """stdout carries the one JSON line only: whatever libraries write to file descriptor 1 (NCCL prints its version
    there) is sent to stderr, the line itself goes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.lines, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None, reasons=sorted(reasons),
                    samples=len(sm))


# ------------------------------------------------------------------------------------------
def prepare_kernels(cfg):
    """PSFs -> kernel1/kernel2 through the host mirror of MVDeconFFT.init (device convolutions)."""
    import spim_registration_b200 as m
    from spim_registration_b200 import synthetic
    psfs = synthetic.make_psfs(cfg["views"], (cfg["k"],) * 3, cfg["extracted"])
    inp = m.MVDeconInput()
    dummy = np.ones((2, 2, 2), np.float32)
    for p in psfs:
        inp.add(m.MVDeconFFT(dummy, 1.0, p, deviceList=(local_device(),)))
    inp.init(cfg["psftype"], 8)
    return psfs, [v.getKernel1() for v in inp.getViews()], [v.getKernel2() for v in inp.getViews()]


def local_device():
    return int(os.environ.get("LOCAL_RANK", "0"))


def device_convolve(lib, t_src, kernel, mode, value=0.0):
    import torch
    from spim_registration_b200 import _lib as L
    out = torch.empty_like(t_src)
    k = np.ascontiguousarray(kernel, dtype=np.float32)
    rc = lib.mvdConvolveDevice(t_src.device.index, C.c_void_p(t_src.data_ptr()), L.int3(tuple(t_src.shape)[::-1]),
                               k.ctypes.data_as(L.c_float_p), L.int3(k.shape[::-1]), {"zero": 0, "const": 1, "mirror": 2}[mode],
                               float(value), C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    L.check(rc, "mvdConvolveDevice")
    return out


def gen_dataset_device(cfg, lib, psfs, sub=None, seed_offset=0):
    """Synthetic views generated on the device (not timed, not the product): sparse sources blurred
    into blobs/beads, blurred by every view's PSF, Poisson-like noise, a no-data slab per view,
    cosine blending weights normalised over the views.  `sub` = (offset_zyx, dim_zyx) of a brick."""
    import torch
    from spim_registration_b200 import synthetic
    dev = torch.device("cuda", local_device())
    Z, Y, X = cfg["shape"]
    if sub is not None and float(Z) * Y * X > 2.0e9:
        # a volume that does not fit one GPU (cfg5): every rank generates its own brick as a small volume of the same
        # statistics (sources, blur, noise); the bricks are not continuous across their faces, which a throughput
        # measurement does not need
        (oz, oy, ox), (dz, dy, dx) = sub
        local = dict(cfg, shape=(dz, dy, dx))
        imgs, _ = gen_dataset_device(dict(local, dense_weights=False), lib, psfs, None, seed_offset=1 + ox + 7 * oy + 13 * oz)
        weights = virtual_weights_device(cfg, lib, sub) if cfg["dense_weights"] else [None] * cfg["views"]
        return imgs, weights
    g = torch.Generator(device=dev)
    g.manual_seed(synthetic.SEED0 + 3 + seed_offset)
    dens = 1.0 / 4096 if cfg["kind"] == "blob" else 1.0 / 32768
    src = (torch.rand((Z, Y, X), generator=g, device=dev) < dens).float() * (0.2 + 0.8 * torch.rand((Z, Y, X), generator=g, device=dev))
    sig = 2.5 if cfg["kind"] == "blob" else 0.8
    kb = synthetic.gaussian_psf((15, 15, 15), sig, sig, 0.0)
    truth = device_convolve(lib, src, kb, "mirror") + 0.01
    del src
    imgs, weights = [], []
    V = cfg["views"]
    for v in range(V):
        k = (psfs[v].astype(np.float64) / psfs[v].astype(np.float64).sum()).astype(np.float32)
        b = device_convolve(lib, truth, k, "mirror")
        g.manual_seed(synthetic.SEED0 + 17 * (v + 1))
        b = b + torch.sqrt(torch.clamp(b, min=0) / 200.0) * torch.randn(b.shape, generator=g, device=dev)
        b.clamp_(min=1e-4)
        b /= b.max()
        w = int(round(0.12 * X))
        if v % 2 == 0:
            b[:, :, :w] = 0
        else:
            b[:, :, X - w:] = 0
        imgs.append(b)
    del truth
    if sub is not None:
        (oz, oy, ox), (dz, dy, dx) = sub
        imgs = [i[oz:oz + dz, oy:oy + dy, ox:ox + dx].contiguous() for i in imgs]
    weights = virtual_weights_device(cfg, lib, sub) if cfg["dense_weights"] else [None] * V
    torch.cuda.synchronize()
    return imgs, weights


def virtual_weights_device(cfg, lib, sub=None):
    """The views' blending weights, evaluated on the device from their registrations like the reference's
    default weight type (virtual weights: ProcessForDeconvolution.java:204-208, NormalizingRandomAccess.java:36-45;
    blending border 8/8/3 and range 12 as in EfficientBayesianBased.java:75-76).  Synthetic registrations: every
    view covers the volume, shifted along x by a view-dependent amount.  `sub` = (offset_zyx, dim_zyx) of a brick.
    Returns device tensors [z,y,x]; on the e2e path this runs inside the timed region instead of an upload."""
    import torch
    from spim_registration_b200 import _lib as L
    dev = torch.device("cuda", local_device())
    Z, Y, X = cfg["shape"]
    V = cfg["views"]
    (oz, oy, ox), (dz, dy, dx) = sub if sub is not None else ((0, 0, 0), (Z, Y, X))
    src_dims = (C.c_int * (3 * V))(*([X, Y, Z] * V))
    inv = np.zeros((V, 3, 4), np.float64)
    for v in range(V):
        inv[v, :, :3] = np.eye(3)
        inv[v, 0, 3] = -(v - (V - 1) / 2.0) * 0.05 * X
    border = (C.c_float * (3 * V))(*([8.0, 8.0, 3.0] * V))
    rng = (C.c_float * (3 * V))(*([12.0, 12.0, 12.0] * V))
    off, dim = L.int3((ox, oy, oz)), L.int3((dx, dy, dz))
    sum_w = torch.empty((dz, dy, dx), dtype=torch.float32, device=dev)
    mn, avg = C.c_int(), C.c_double()
    L.check(lib.mvdVirtualWeightSum(dev.index, V, src_dims, inv.ctypes.data_as(L.c_double_p), border, rng, off, dim,
                                    C.c_void_p(sum_w.data_ptr()), C.byref(mn), C.byref(avg), 1), "mvdVirtualWeightSum")
    weights = []
    for v in range(V):
        w = torch.empty((dz, dy, dx), dtype=torch.float32, device=dev)
        L.check(lib.mvdVirtualWeight(dev.index, L.int3((X, Y, Z)), inv[v].ctypes.data_as(L.c_double_p), (C.c_float * 3)(8.0, 8.0, 3.0),
                                     (C.c_float * 3)(12.0, 12.0, 12.0), off, dim, C.c_void_p(sum_w.data_ptr()), 1.0,
                                     C.c_void_p(w.data_ptr()), 1), "mvdVirtualWeight")
        weights.append(w)
    return weights


# ------------------------------------------------------------------------------------------
def multi_gpu_parity(lib, dist, rank, world, dev_index, iters=3):
    """world > 1: the sharded path (bricks + ghost layers, fused peer-store exchange over CUDA IPC with the barrier
    between steps, and the NCCL send/recv exchange) against a single-context run of the same small workload on
    rank 0, before anything is timed.  cfg0 (3 views, 192x160x96, PSF 15^3: view 0 runs the direct separable
    kernels, views 1-2 the FFT engine).  Returns the `parity_multi_gpu` record; every rank gets the same verdict."""
    import torch
    from spim_registration_b200 import _lib as L
    from spim_registration_b200.parallel import ShardedDeconvolution, brick_grid, brick_of
    cfg = CONFIGS[0]
    Z, Y, X = cfg["shape"]
    dim_xyz = (X, Y, Z)
    psfs, k1, k2 = prepare_kernels(cfg)
    imgs, weights = gen_dataset_device(cfg, lib, psfs, None)      # the same seeded volume on every rank
    kmax_xyz = [max(max(k.shape[::-1][d] for k in k1), max(k.shape[::-1][d] for k in k2)) for d in range(3)]
    grid = brick_grid(world, dim_xyz, kmax_xyz)
    off, size, _ = brick_of(rank, grid, dim_xyz)
    sl = tuple(slice(off[d], off[d] + size[d]) for d in (2, 1, 0))
    b_imgs = [i[sl].contiguous() for i in imgs]
    b_w = [w[sl].contiguous() for w in weights]
    dev = torch.device("cuda", dev_index)
    full = {}
    for mode in ("fused", "nccl"):
        dec = ShardedDeconvolution(dim_xyz, b_imgs, b_w, k1, k2, LAMBDA, dist, dev_index, grid=grid)
        dec.init_psi_average()
        if mode == "fused":
            dec.attach_peers()
        else:
            dec._exchange(0)
        dec.run_iterations(iters)
        dec.mask_never_updated()
        vol = torch.zeros((Z, Y, X), dtype=torch.float32, device=dev)
        vol[sl] = torch.from_numpy(dec.get_psi_brick()).to(dev)
        dist.all_reduce(vol)                                       # bricks are disjoint: the sum assembles the volume
        full[mode] = vol
        dec.close()
    verdict = torch.zeros(4, dtype=torch.float64, device=dev)
    if rank == 0:
        one = ShardedDeconvolution(dim_xyz, imgs, weights, k1, k2, LAMBDA, None, dev_index)
        one.init_psi_average()
        one.run_iterations(iters)
        one.mask_never_updated()
        ref = torch.from_numpy(one.get_psi_brick()).to(dev).double()
        one.close()
        verdict[0] = float((full["fused"].double() - ref).norm() / ref.norm())
        verdict[1] = float((full["nccl"].double() - ref).norm() / ref.norm())
        verdict[2] = 1.0 if torch.equal(full["fused"], full["nccl"]) else 0.0
        verdict[3] = float((full["fused"].double() - full["nccl"].double()).norm() / ref.norm())
    dist.broadcast(verdict, 0)
    r = dict(workload=cfg["name"], iterations=iters, brick_grid_xyz=list(grid), rel_l2=float(verdict[0]), rel_l2_nccl_exchange=float(verdict[1]),
             bit_identical=bool(verdict[2] == 1.0), rel_l2_fused_vs_nccl=float(verdict[3]),
             compared="fused peer-store exchange (CUDA IPC + barrier) and NCCL exchange vs a single-context run on rank 0; the single context "
             "tiles the volume differently, so rel_l2 is at fp32 rounding level, not 0; bit_identical = fused vs NCCL exchange: the fused "
             "epilogue also fills the float4 padding columns of the x ghost layers (never consumed by the kernel support, but inside the "
             "FFT windows, i.e. they change rounding noise), the NCCL exchange leaves them at their initial value",
             tolerance=1e-4, tolerance_fused_vs_nccl=1e-6)
    r["ok"] = bool(r["rel_l2"] <= 1e-4 and r["rel_l2_nccl_exchange"] <= 1e-4 and r["rel_l2_fused_vs_nccl"] <= 1e-6)
    del imgs, weights, b_imgs, b_w, full
    torch.cuda.empty_cache()
    return r


# ------------------------------------------------------------------------------------------
def run_ours(args, cfg):
    import torch
    import torch.distributed as dist
    import spim_registration_b200 as m
    from spim_registration_b200 import _lib as L
    from spim_registration_b200.parallel import ShardedDeconvolution, brick_grid, brick_of

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dev_index = local_device()
    torch.cuda.set_device(dev_index)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev_index))
    lib = m.load()
    parity = None
    Z, Y, X = cfg["shape"]
    dim_xyz = (X, Y, Z)
    V = cfg["views"]
    nvox = float(X) * Y * Z
    psfs, k1, k2 = prepare_kernels(cfg)
    kmax_xyz = [max(max(k.shape[::-1][d] for k in k1), max(k.shape[::-1][d] for k in k2)) for d in range(3)]
    grid = brick_grid(world, dim_xyz, kmax_xyz)
    off, size, _ = brick_of(rank, grid, dim_xyz)
    sub = None if world == 1 else (off[::-1], size[::-1])
    imgs, weights = gen_dataset_device(cfg, lib, psfs, sub)
    dz = dist if world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident throughput (`value`) ------------------------------------------------
    fused = world > 1 and os.environ.get("MVD_HALO", "fused") == "fused"
    dec = ShardedDeconvolution(dim_xyz, imgs, weights, k1, k2, LAMBDA, dz, dev_index, grid=grid)
    dec.init_psi_average()
    if fused:
        dec.attach_peers()  # step kernels store ghost voxels into the neighbours' arrays over NVLink
    elif world > 1:
        dec._exchange(0)
    for _ in range(args.warmup):
        dec.run_iterations(1)
    barrier()
    sampler = ClockSampler(dev_index)
    if rank == 0:
        sampler.start()
    launches0 = dec.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()  # `ncu --profile-from-start off` captures exactly the timed steps
    e0.record()
    dec.run_iterations(args.steps)
    e1.record()
    barrier()
    torch.cuda.profiler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    launches = dec.kernel_launches() - launches0
    # per-kernel durations: the same steps again with every launch bracketed by CUDA events on one
    # stream (the timed region above deals independent tile batches to several streams, where a
    # launch's elapsed time would include its neighbours)
    prof_steps = min(args.steps, 2)
    L.check(lib.mvdProfileStart(dec.ctx), "mvdProfileStart")
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    dec.run_iterations(prof_steps)
    p1.record()
    barrier()
    prof_ms = p0.elapsed_time(p1)
    pms, pl, pb = (C.c_double * 3)(), (C.c_longlong * 3)(), (C.c_double * 3)()
    L.check(lib.mvdProfileStop(dec.ctx, pms, pl, pb), "mvdProfileStop")
    dec.close()
    del dec
    value = nvox * args.steps / (ms * 1e-3) / 1e9

    # ---- end to end through the public API with host buffers (`e2e`) ---------------------------
    h_imgs = [torch.empty(i.shape, dtype=torch.float32).pin_memory() for i in imgs]
    for h, d in zip(h_imgs, imgs):
        h.copy_(d)
    dense = cfg["dense_weights"]
    del imgs, weights
    torch.cuda.empty_cache()
    h_psi = torch.empty(size[::-1], dtype=torch.float32).pin_memory()
    np_imgs = [h.numpy() for h in h_imgs]
    h2d = sum(a.nbytes for a in np_imgs)
    d2h = h_psi.numpy().nbytes
    def e2e_pass():
        barrier()
        t0 = time.perf_counter()
        # the reference's default weight type: no weight images exist on the host, the weights follow from the
        # views' registrations -- evaluated on the device here, inside the timed region
        e2e_w = virtual_weights_device(cfg, lib, sub) if dense else [1.0 / V] * V
        torch.cuda.synchronize()
        tw = time.perf_counter()
        # H2D of every view, asynchronous (MVD_MEM_HOST_ASYNC): view v's first step waits for its own upload on the device, so the
        # loop starts behind view 0 while the other views are still crossing PCIe; the pinned buffers live until the end
        dec = ShardedDeconvolution(dim_xyz, np_imgs, e2e_w, k1, k2, LAMBDA, dz, dev_index, grid=grid, async_upload=True)
        del e2e_w
        t1 = time.perf_counter()
        # the neighbours' arrays are opened (CUDA IPC) while the views are still uploading; psi init waits for the uploads on the device
        if fused:
            dec.attach_peers()
        dec.init_psi_average()
        if world > 1 and not fused:
            dec._exchange(0)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        dec.run_iterations(args.steps)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        dec.mask_never_updated()
        dec.get_psi_brick(h_psi.numpy())  # D2H of the result
        barrier()
        t4 = time.perf_counter()
        e2e_phases = {"virtual_weights_on_device": tw - t0, "create_context_enqueue_h2d_views_spectra": t1 - tw, "attach_peers_wait_uploads_psi_init": t2 - t1, "iterations": t3 - t2, "mask_d2h": t4 - t3}
        e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        e2e_s = float(e2e_s.item())
        dec.close()
        return e2e_s, e2e_phases
    # one untimed pass first (--e2e-warmup), like the loop's warm-up steps: the first pass of this leg in a process took 3.3 s on one
    # fresh box (0.5 s in the blending-weight kernels, 1.8 s in context creation) and 1.3 - 1.4 s on the others; the warm-up pass's
    # time is in the JSON line beside the timed one
    e2e_warm = [e2e_pass()[0] for _ in range(args.e2e_warmup)]
    e2e_s, e2e_phases = e2e_pass()
    e2e_value = nvox * args.steps / e2e_s / 1e9

    # ---- multi-GPU parity leg: after the timed regions (its NCCL point-to-point exchange leaves communicator state behind that
    # slowed a timed region run right after it by ~5 %), before anything is printed: a difference fails the run
    if world > 1 and not args.no_parity:
        del h_imgs, np_imgs, h_psi
        torch.cuda.empty_cache()
        parity = multi_gpu_parity(lib, dist, rank, world, dev_index)
        if not parity["ok"]:
            if rank == 0:
                emit({"error": "multi-GPU result differs from the single-GPU result", "parity_multi_gpu": parity, "n_gpus": world})
            dist.destroy_process_group()
            sys.exit(3)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    names = ("fwd_xy", "z_pass", "inv_xy")
    kern = {names[i]: dict(ms=pms[i], launches=int(pl[i]), alg_gb=pb[i] / 1e9,
                           gbs=(pb[i] / 1e9) / (pms[i] * 1e-3) if pms[i] > 0 else 0.0) for i in range(3)}
    top = max(range(3), key=lambda i: pms[i])
    achieved = kern[names[top]]["gbs"]
    # DRAM bytes per launch of the dominant kernel: measured once per round with `ncu --set full`
    # (dram__bytes_read.sum + dram__bytes_write.sum over complete convolutions) as a ratio to the byte
    # model, profiles/*_dram_traffic.json; applied to this run's algorithmic bytes per launch
    traffic, traffic_src = None, None
    import glob
    tjs = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_dram_traffic.json")), key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])
    tj = tjs[-1] if tjs else ""
    if tj and pl[top] > 0:
        tk = json.load(open(tj))["kernels"]
        ratios = [v["traffic_over_model"] for k, v in tk.items() if k.startswith(names[top])]
        if ratios:
            traffic = sum(ratios) / len(ratios) * pb[top] / pl[top]
            traffic_src = "profiles/%s (ncu --set full, dram bytes / byte model = %.3f) x this run's algorithmic bytes per launch" % (os.path.basename(tj), sum(ratios) / len(ratios))
    loop_bytes = ALG_BYTES_PER_VOXEL_VIEW[cfg["dense_weights"]] * V * nvox * args.steps / world
    line = {
        "metric": "Gvoxel-iterations/sec (all views)", "value": value, "unit": "Gvoxel-iterations/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["name"], "views": V, "volume_xyz": list(dim_xyz), "psf": cfg["k"], "iteration_type": cfg["psftype"],
                   "lambda": LAMBDA, "step": "one iteration over all views", "brick_grid_xyz": list(grid),
                   "halo": "none" if world == 1 else ("fused: epilogue stores into peer ghost layers (NVLink) + barrier" if fused else "NCCL send/recv per axis"),
                   "l2": "inputs larger than L2 (%.1f GB resident per GPU)" % ((2 * V + 2) * nvox * 4 / world / 1e9),
                   "scratch_mb": int(os.environ.get("MVD_SCRATCH_MB", "4096")), "batch_streams": "1 + %d side lanes for sub-wave tile batches" % (int(os.environ.get("MVD_LANES", "4")) - 1)},
        "roofline": {"bound": "hbm", "kernel": names[top], "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "achieved_bytes_per_launch": pb[top] / pl[top] if pl[top] else None, "peak_source": peak_src,
                     "share_of_step": (pms[top] / prof_steps) / (ms / args.steps), "kernels": kern,
                     "timing": "CUDA events around every launch, %d serial step(s) after the timed region: kernels %.1f ms/step, wall %.1f ms/step with the event calls; timed region %.1f ms/step (MVD_LANES=%s); share_of_step = kernel ms per step / timed ms per step (views on the direct separable path are not in `kernels`)"
                               % (prof_steps, sum(pms) / prof_steps, prof_ms / prof_steps, ms / args.steps, os.environ.get("MVD_LANES", "4")),
                     "loop": {"alg_bytes_per_voxel_view_iteration": ALG_BYTES_PER_VOXEL_VIEW[cfg["dense_weights"]],
                              "achieved": loop_bytes / 1e9 / (ms * 1e-3), "frac": loop_bytes / 1e9 / (ms * 1e-3) / peak,
                              "frac_of_8TBs": loop_bytes / 1e9 / (ms * 1e-3) / 8000.0}},
        "e2e": {"value": e2e_value, "unit": "Gvoxel-iterations/s", "h2d_bytes_per_step": h2d / args.steps,
                "d2h_bytes_per_step": d2h / args.steps, "seconds": e2e_s, "phase_seconds_rank0": e2e_phases,
                "untimed_warmup_pass_seconds": e2e_warm,
                "includes": "blending weights evaluated on the device from the views' registrations (the reference's default 'virtual weights'; kept resident), context creation, H2D of all views (pinned, on a copy stream: overlaps the first steps), kernel spectra, psi init, %d iterations, mask, D2H of psi" % args.steps},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if parity is not None:
        line["parity_multi_gpu"] = parity
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_sample(cfg, k1, k2)
    emit(line)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------
def run_single_process(args, cfg):
    """--single-process: N GPUs of the box driven from ONE process through the multi-device handle of the C ABI
    (mvdCreateMulti: worker thread per device, event-based neighbour barrier, no torch.distributed / NCCL) -- how
    the reference's Java host would use several devices (MVDeconFFT.java:426-445).  Host buffers in, host buffers out."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    import spim_registration_b200 as m
    from spim_registration_b200 import _lib as L
    lib = m.load()
    n = args.gpus
    torch.cuda.set_device(0)
    Z, Y, X = cfg["shape"]
    V = cfg["views"]
    nvox = float(X) * Y * Z
    psfs, k1, k2 = prepare_kernels(cfg)
    imgs, weights = gen_dataset_device(cfg, lib, psfs, None)
    h_imgs = [i.cpu().numpy() for i in imgs]
    h_w = [w.cpu().numpy() if w is not None else None for w in weights]
    del imgs, weights
    torch.cuda.empty_cache()
    kmax = [max(max(k.shape[::-1][d] for k in k1), max(k.shape[::-1][d] for k in k2)) for d in range(3)]
    t0 = time.perf_counter()
    h = C.c_void_p()
    L.check(lib.mvdCreateMulti(C.byref(h), (C.c_int * n)(*range(n)), n, L.int3((X, Y, Z)), L.int3(kmax), LAMBDA, 1e-4), "mvdCreateMulti")
    for v in range(V):
        wp, wc = (h_w[v].ctypes.data_as(C.c_void_p), 1.0) if h_w[v] is not None else (None, 1.0 / V)
        L.check(lib.mvdMultiAddView(h, h_imgs[v].ctypes.data_as(C.c_void_p), wp, wc, k1[v].ctypes.data_as(L.c_float_p), L.int3(k1[v].shape[::-1]),
                                    k2[v].ctypes.data_as(L.c_float_p), L.int3(k2[v].shape[::-1])), "mvdMultiAddView")
    avg = C.c_double()
    L.check(lib.mvdMultiInitPsiAverage(h, C.byref(avg)), "mvdMultiInitPsiAverage")
    t1 = time.perf_counter()
    L.check(lib.mvdMultiRunIterations(h, args.warmup, None, None), "mvdMultiRunIterations")
    launches0 = lib.mvdKernelLaunchCount()
    t2 = time.perf_counter()
    L.check(lib.mvdMultiRunIterations(h, args.steps, None, None), "mvdMultiRunIterations")   # returns when every device has finished
    t3 = time.perf_counter()
    launches = lib.mvdKernelLaunchCount() - launches0
    L.check(lib.mvdMultiMaskNeverUpdated(h), "mvdMultiMaskNeverUpdated")
    psi = np.empty((Z, Y, X), np.float32)
    L.check(lib.mvdMultiGetPsi(h, psi.ctypes.data_as(C.c_void_p)), "mvdMultiGetPsi")
    t4 = time.perf_counter()
    lib.mvdDestroyMulti(h)
    ms = (t3 - t2) * 1e3
    e2e_s = (t1 - t0) + (t3 - t2) + (t4 - t3)
    emit({"metric": "Gvoxel-iterations/sec (all views)", "value": nvox * args.steps / (ms * 1e-3) / 1e9, "unit": "Gvoxel-iterations/s", "n_gpus": n,
          "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
          "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "single-process C ABI (mvdCreateMulti)",
          "config": {"workload": cfg["name"], "views": V, "volume_xyz": [X, Y, Z], "psf": cfg["k"], "iteration_type": cfg["psftype"], "lambda": LAMBDA,
                     "step": "one iteration over all views", "halo": "epilogue stores into peer ghost layers (NVLink) + per-step neighbour events",
                     "timing": "host clock around mvdMultiRunIterations, which returns after every device's stream has drained"},
          "e2e": {"value": nvox * args.steps / e2e_s / 1e9, "unit": "Gvoxel-iterations/s", "h2d_bytes_per_step": sum(a.nbytes for a in h_imgs + [w for w in h_w if w is not None]) / args.steps,
                  "d2h_bytes_per_step": psi.nbytes / args.steps, "seconds": e2e_s, "includes": "context creation, H2D of views and weights from pageable host arrays, spectra, psi init, iterations, mask, D2H"},
          "gpu_launches": int(launches), "psi_mean": float(psi.mean(dtype=np.float64))})


# ------------------------------------------------------------------------------------------
def reference_crop(cfg, blocks_xyz=(2, 2, 1)):
    """The bounded sample of the workload the CPU arm runs: a crop of blocks_xyz effective regions of the reference's
    256^3 blocks (64^3 for tiny configs) -- 452 x 452 x 226 voxels at K = 31 -- the same at every N and in every step."""
    Z, Y, X = cfg["shape"]
    block = [256, 256, 256] if min(cfg["shape"]) >= 256 else [64, 64, 64]
    e = block[0] - cfg["k"] + 1
    crop = [min(n, e * b) for n, b in zip((Z, Y, X), blocks_xyz[::-1])]
    return block, crop


def cpu_sample(cfg, k1, k2, budget_s=20.0, steps=1, crop=None):
    """CPU port of the reference's CPU path (oracle/cpu_baseline.py) on a bounded sample: a fixed crop of the workload
    (reference_crop) in the reference's blocks, all views, `steps` iterations; throughput per voxel.  numpy-generated
    inputs of the same statistics.  `crop` = (z, y, x) overrides the sample (the whole volume for --full-iteration)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_baseline as cb
    V = cfg["views"]
    cores = cb.host_threads()
    block, default_crop = reference_crop(cfg)
    crop = list(crop or default_crop)
    rng = np.random.default_rng(7)
    imgs = [(rng.random(crop, dtype=np.float32) * 0.9 + 0.05) for _ in range(V)]
    w = [np.full(crop, 1.0 / V, np.float32) if cfg["dense_weights"] else np.float32(1.0 / V) for _ in range(V)]
    psi = np.full(crop, 0.3, np.float32)
    _, t = cb.run_iterations(imgs, w, k1, k2, steps, LAMBDA, block, psi, cores)
    vox = float(np.prod(crop)) * steps
    return {"value": vox / t / 1e9, "unit": "Gvoxel-iterations/s", "cores": cores, "kind": "port",
            "sample": "%d view(s), crop %dx%dx%d (xyz) of the workload in %d^3 blocks, %d iteration(s), %.1f s, scipy pocketfft float32 with %d workers"
                      % (V, crop[2], crop[1], crop[0], block[0], steps, t, cores), "seconds": t}


def run_reference(args, cfg):
    """--impl reference: the CPU port on the host cores (rank 0 only; the other ranks leave at once).  Every step runs
    the same fixed crop with all host threads -- thread count and affinity are set here, so that the figure does not
    depend on how the process was launched (torchrun exports OMP_NUM_THREADS=1 to its workers)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    ncpu = os.cpu_count() or 1
    try:
        os.sched_setaffinity(0, range(ncpu))
    except (AttributeError, OSError):
        pass
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(ncpu)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_np as o
    from spim_registration_b200 import synthetic
    psfs = synthetic.make_psfs(cfg["views"], (cfg["k"],) * 3, cfg["extracted"])
    k1, k2 = o.init_kernels(psfs, cfg["psftype"], 8)
    vals, secs, sample = [], [], None
    for i in range(args.warmup + args.steps):
        r = cpu_sample(cfg, k1, k2)
        if i >= args.warmup:
            vals.append(r["value"])
            secs.append(r["seconds"])
        sample = r
    value = float(np.mean(vals))
    Z, Y, X = cfg["shape"]
    full = None
    if args.full_iteration:   # one iteration of the WHOLE workload, once, beside the crop figure
        r = cpu_sample(cfg, k1, k2, crop=(Z, Y, X))
        full = {"value": r["value"], "seconds": r["seconds"], "sample": r["sample"]}
    line = {"impl": "reference", "metric": "Gvoxel-iterations/sec (all views)", "value": value, "unit": "Gvoxel-iterations/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": float(np.mean(secs)) * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": cfg["name"], "views": cfg["views"], "volume_xyz": [X, Y, Z], "psf": cfg["k"],
                       "iteration_type": cfg["psftype"], "lambda": LAMBDA,
                       "step": "one iteration over all views on a fixed crop of the workload (see cpu_baseline.sample); value = crop voxels / time, ms_per_step = the time of the crop",
                       "spread": {"min": float(np.min(vals)), "max": float(np.max(vals))}},
            "cpu_baseline": {"value": value, "unit": "Gvoxel-iterations/s", "cores": sample["cores"], "kind": "port", "sample": sample["sample"]},
            "e2e": {"value": value, "unit": "Gvoxel-iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if full:
        line["full_iteration"] = full
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--single-process", action="store_true", help="drive --gpus N devices from one process through mvdCreateMulti")
    ap.add_argument("--no-parity", action="store_true", help="skip the multi-GPU parity leg (world > 1)")
    ap.add_argument("--cpu-seconds", type=float, default=20.0)
    ap.add_argument("--e2e-warmup", type=int, default=1, help="untimed passes of the end-to-end leg before the timed one")
    ap.add_argument("--full-iteration", action="store_true", help="--impl reference: also time one iteration of the whole workload")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    quiet_stdout()
    if cfg.get("min_gpus", 1) > max(args.gpus, int(os.environ.get("WORLD_SIZE", "1"))) and args.impl == "ours":
        if int(os.environ.get("RANK", "0")) == 0:
            emit({"error": "%s needs %d GPUs (%.0f GB resident)" % (cfg["name"], cfg["min_gpus"], (2 * cfg["views"] + 2) * float(np.prod(cfg["shape"])) * 4 / 1e9)})
        sys.exit(2)
    if args.impl == "reference":
        run_reference(args, cfg)
    elif args.single_process:
        run_single_process(args, cfg)
    else:
        run_ours(args, cfg)


if __name__ == "__main__":
    main()
