"""Timed CPU baseline ("port") of the reference's CPU path -- TEST/BENCH INFRASTRUCTURE ONLY.

The reference's Java CPU path cannot run here (no JVM).  This restates its *structure*
 with the fastest FFT available on the box, scipy's multithreaded
pocketfft in float32, so the baseline is if anything faster than the Java original (Mines JTK
through per-element accessors, J/fftMethods/FFTMethods.java:934-992):

  * block-wise copyBlock -> float32 R2C FFT -> pointwise multiply with the cached kernel FFT ->
    inverse FFT -> pasteBlock (MVDeconFFT.java:363-535, MVDeconFFTThreads.java:18-49,
    Block.java:114-195, J/fftMethods/FFTConvolution.java:470-551);
  * the quotient and the update as separate elementwise passes, split over a pool of host threads
    like the reference's portions (MVDeconvolution.java:446-463, 473-525, 582-703);
  * psi updated after every view (MVDeconvolution.java:353-441).

Only bench.py's cpu_baseline leg / --impl reference and tests may import this.
"""
import os
import time

import numpy as np
from scipy import fft as sfft

import oracle_np as o

F32 = np.float32


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class BlockConvolver:
    """FFTConvolution on fixed-size blocks with the kernel FFT kept (keepImgFFT=false,
    kernel FFT cached: MVDeconFFT.java:373-379)."""

    def __init__(self, kernel, block_size_xyz, workers):
        self.k = np.asarray(kernel, dtype=F32)
        self.bs = list(block_size_xyz)
        self.workers = workers
        shape = tuple(self.bs[::-1])
        kc = np.zeros(shape, dtype=F32)
        K = self.k.shape
        # kernel centre to index 0, periodic (FFTConvolution.java:506-516)
        idx = [np.mod(np.arange(K[a]) - K[a] // 2, shape[a]) for a in range(3)]
        kc[np.ix_(*idx)] = self.k
        self.kf = sfft.rfftn(kc, workers=workers)
        self.shape = shape

    def __call__(self, block):
        f = sfft.rfftn(block, workers=self.workers)
        f *= self.kf
        return sfft.irfftn(f, s=self.shape, workers=self.workers).astype(F32, copy=False)


def convolve_blocks(src, conv, mode, value, blocks):
    out = np.empty_like(src)
    for b in blocks:
        data = o.copy_block(src, b, mode, value)
        o.paste_block(out, conv(data), b)
    return out


def _slabs(n, parts):
    edges = np.linspace(0, n, min(parts, n) + 1).astype(int)
    return [(int(a), int(b)) for a, b in zip(edges[:-1], edges[1:]) if b > a]


def threaded_quotient(pool, parts, blurred, img):
    """computeQuotient on z-slabs, one per thread (the reference splits the image into portions for a
    thread pool, MVDeconvolution.java:446-463, 473-525); numpy releases the GIL inside the array operations."""
    out = np.empty_like(blurred)

    def work(ab):
        a, b = ab
        out[a:b] = o.compute_quotient(blurred[a:b], img[a:b])
    list(pool.map(work, _slabs(blurred.shape[0], parts)))
    return out


def threaded_final_values(pool, parts, psi, integral, weight, lam):
    """computeFinalValues on z-slabs (MVDeconvolution.java:582-659)."""
    out = np.empty_like(psi)
    dense = not np.isscalar(weight) and np.ndim(weight) > 0

    def work(ab):
        a, b = ab
        out[a:b], s, mx = o.compute_final_values(psi[a:b], integral[a:b], weight[a:b] if dense else weight, lam)
        return s, mx
    stats = list(pool.map(work, _slabs(psi.shape[0], parts)))
    return out, sum(s for s, _ in stats), max(mx for _, mx in stats)


def run_iterations(imgs, weights, k1, k2, iterations, lam, block_size_xyz, psi, workers=None):
    """`iterations` passes over all views on the given volume; returns (psi, seconds)."""
    workers = workers or host_threads()
    shape_xyz = list(np.asarray(imgs[0]).shape[::-1])
    V = len(imgs)
    conv1 = [BlockConvolver(k1[v], block_size_xyz, workers) for v in range(V)]
    conv2 = [BlockConvolver(k2[v], block_size_xyz, workers) for v in range(V)]
    blocks1 = [o.divide_into_blocks(shape_xyz, list(k1[v].shape[::-1]), block_size_xyz) for v in range(V)]
    blocks2 = [o.divide_into_blocks(shape_xyz, list(k2[v].shape[::-1]), block_size_xyz) for v in range(V)]
    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=workers) as pool:
        for _ in range(iterations):
            for v in range(V):
                t1 = convolve_blocks(psi, conv1[v], "mirror", 0.0, blocks1[v])
                t1 = threaded_quotient(pool, workers, t1, imgs[v])
                t2 = convolve_blocks(t1, conv2[v], "const", 1.0, blocks2[v])
                psi, _, _ = threaded_final_values(pool, workers, psi, t2, weights[v], lam)
    return psi, time.perf_counter() - t0
