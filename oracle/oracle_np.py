"""CPU oracle (numpy/scipy) of the multi-view deconvolution hot path -- TEST INFRASTRUCTURE ONLY.

This file restates, function by function, the algorithm of PreibischLab/SPIM_Registration's
multi-view Richardson-Lucy loop so that the CUDA library can be checked against it.  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import it; the product
(spim_registration_b200/) never does.

PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path and
cannot be executed here (no JVM, dependencies not vendored).  The oracle
is pinned only against the known-answer values derivable from the reference's own `main`
methods (Block.java:367-402 tiling, MVDeconvolution.java:707-714 tikhonov table) -- see
tests/test_oracle.py.

Array convention: numpy arrays are indexed [z, y, x] (C order), i.e. x fastest, matching the
reference's ArrayImg layout index = x + X*(y + Y*z).  Citations are relative to
/root/reference/src/main/java/ (abbreviated J/).

Convolutions are evaluated exactly (float64 FFT on the explicitly extended array, then cast to
float32 wherever the reference stores a FloatType) instead of through a float32 FFT as the
reference does (J/fftMethods/FFTConvolution.java:470-551); the convention is TRUE convolution
with the kernel centre at floor(K/2) (FFTConvolution.java:506-516), see SURVEY.md 8c caveat.
"""
import math
import os

import numpy as np
from scipy import fft as sfft
from scipy import signal

WORKERS = int(os.environ.get("ORACLE_WORKERS", "0")) or (os.cpu_count() or 1)

F32 = np.float32
MIN_VALUE = F32(0.0001)  # MVDeconvolution.java:49 / MVDeconInput.java:12

PSFTYPES = ("OPTIMIZATION_II", "OPTIMIZATION_I", "EFFICIENT_BAYESIAN", "INDEPENDENT")  # MVDeconFFT.java:28


# --------------------------------------------------------------------------------------------
# J/spim/process/fusion/FusionHelper.java:236-260
def divide_into_portions(image_size, num_portions):
    chunk = image_size // num_portions
    mod = image_size % num_portions
    out = []
    for pid in range(num_portions):
        start = pid * chunk
        loop = chunk + mod if pid == num_portions - 1 else chunk
        out.append((start, loop))
    return out


# J/spim/process/fusion/deconvolution/AdjustInput.java:44-100
def sum_img(img, num_threads):
    """sumImg including its quirk: the partial sum of the portion that started first
    (deterministically restated as portion 0) is added twice (AdjustInput.java:93-97)."""
    flat = np.asarray(img, dtype=F32).reshape(-1).astype(np.float64)
    portions = divide_into_portions(flat.size, num_threads * 2)
    sums = [math.fsum(flat[s:s + n]) for (s, n) in portions]
    return math.fsum([sums[0]] + sums)


# AdjustInput.java:29-36
def norm_img(img, num_threads):
    s = sum_img(img, num_threads)
    return (np.asarray(img, dtype=F32).astype(np.float64) / s).astype(F32)


# J/spim/process/fusion/deconvolution/Mirror.java:31-108 (sequential semantics)
def mirror(img, dimension):
    """dimension counts x=0, y=1, z=2.  Odd sizes: a true flip.  Even sizes: the middle pair is
    swapped twice and therefore stays (Mirror.java:72, sizeMirrorH = dim/2)."""
    a = np.array(img, dtype=F32, copy=True)
    axis = a.ndim - 1 - dimension
    n = a.shape[axis]
    a = np.moveaxis(a, axis, 0)
    out = a.copy()
    half = n // 2
    for p in range(half + 1):
        q = n - 1 - p
        tmp = out[p].copy()
        out[p] = out[q]
        out[q] = tmp
    return np.ascontiguousarray(np.moveaxis(out, 0, axis))


# MVDeconFFT.java:315-323
def compute_inverted_kernel(kernel):
    k = np.asarray(kernel, dtype=F32)
    for d in range(k.ndim):
        k = mirror(k, d)
    return k


# MVDeconFFT.java:305-313, 325-333 (float multiplication repeated numViews-1 times)
def compute_exponential_kernel(kernel, num_views):
    k = np.asarray(kernel, dtype=F32)
    r = k.copy()
    for _ in range(1, num_views):
        r = (r * k).astype(F32)
    return r


# --------------------------------------------------------------------------------------------
# out-of-bounds rules
def extend(a, halo_lo, halo_hi, mode, value=0.0):
    """Extend array a by halo_lo/halo_hi (per numpy axis) voxels.
    mode 'mirror' = Views.extendMirrorSingle (edge voxel not repeated), 'const' = extendValue,
    'zero' = extendZero."""
    a = np.asarray(a)
    pad = list(zip(halo_lo, halo_hi))
    if mode == "mirror":
        # np.pad 'reflect' needs halo < n; fall back to explicit index maps otherwise
        if all(max(lo, hi) < n for (lo, hi), n in zip(pad, a.shape)):
            return np.pad(a, pad, mode="reflect")
        out = a
        for ax, ((lo, hi), n) in enumerate(zip(pad, a.shape)):
            idx = np.arange(-lo, n + hi)
            if n == 1:
                idx = np.zeros_like(idx)
            else:
                p = 2 * n - 2
                idx = np.mod(idx, p)
                idx = np.where(idx >= n, p - idx, idx)
            out = np.take(out, idx, axis=ax)
        return out
    if mode == "const":
        return np.pad(a, pad, mode="constant", constant_values=value)
    if mode == "zero":
        return np.pad(a, pad, mode="constant", constant_values=0.0)
    raise ValueError(mode)


def convolve_extended(f, kernel, mode, value=0.0):
    """conv(ext(f), k)(x) = sum_j ext(f)(x + K//2 - j) k(j) for x in the domain of f.
    (MVDeconFFT.java:363-535 semantics, FFTConvolution.java:506-516 kernel centre.)  float32 out."""
    k = np.asarray(kernel, dtype=np.float64)
    K = k.shape
    hr = [s // 2 for s in K]           # reads up to x + K//2
    hl = [s - 1 - s // 2 for s in K]   # and down to x - (K-1-K//2)
    fe = extend(np.asarray(f, dtype=np.float64), hl, hr, mode, value)
    with sfft.set_workers(WORKERS):  # lines are independent: same values for every worker count
        out = signal.fftconvolve(fe, k, mode="valid")
    assert out.shape == np.asarray(f).shape
    return out.astype(F32)


def convolve_direct(f, kernel, mode, value=0.0):
    """FFT-free O(N K^3) cross-check of convolve_extended for tiny arrays."""
    k = np.asarray(kernel, dtype=np.float64)
    K = k.shape
    hr = [s // 2 for s in K]
    hl = [s - 1 - s // 2 for s in K]
    fe = extend(np.asarray(f, dtype=np.float64), hl, hr, mode, value)
    out = np.zeros(np.asarray(f).shape, dtype=np.float64)
    Z, Y, X = out.shape
    for jz in range(K[0]):
        for jy in range(K[1]):
            for jx in range(K[2]):
                # ext index = x + hl + (K//2 - j) ... with ext origin at -hl: x + K//2 - j + hl
                oz, oy, ox = hl[0] + hr[0] - jz, hl[1] + hr[1] - jy, hl[2] + hr[2] - jx
                out += k[jz, jy, jx] * fe[oz:oz + Z, oy:oy + Y, ox:ox + X]
    return out.astype(F32)


# --------------------------------------------------------------------------------------------
# kernel preparation, MVDeconFFT.init (MVDeconFFT.java:162-303) driven by MVDeconInput.init
# (MVDeconInput.java:43-49): views are initialised IN ORDER, so while view v is prepared the
# kernel1 of views w < v is already normalised and that of views w > v is not yet.
def separable_convolve(f, kx, ky, kz, mode, value=0.0):
    """The reference's second native hook (CUDASeparableConvolution.java:13-17; implementation in the
    external SeparableConvolutionCUDALib, not in the reference tree -- parity unpinned): three
    in-place 1-D passes x, y, z, each seeing the image extended by `mode` ("zero", "const" = value,
    "clamp" = extend last pixel); out(p) = sum_j k[j] f(p - j + K // 2).  None skips an axis.
    f is [z, y, x]; double accumulation, float32 between the passes like an in-place float image."""
    a = np.asarray(f, dtype=np.float32)
    for axis, k in ((2, kx), (1, ky), (0, kz)):
        if k is None:
            continue
        k = np.asarray(k, dtype=np.float64)
        K = len(k)
        lo, hi = K - 1 - K // 2, K // 2
        pad = [(0, 0)] * 3
        pad[axis] = (lo, hi)
        if mode == "zero":
            e = np.pad(a.astype(np.float64), pad, mode="constant", constant_values=0.0)
        elif mode == "const":
            e = np.pad(a.astype(np.float64), pad, mode="constant", constant_values=float(value))
        elif mode == "clamp":
            e = np.pad(a.astype(np.float64), pad, mode="edge")
        else:
            raise ValueError(mode)
        n = a.shape[axis]
        out = np.zeros(a.shape, np.float64)
        for j in range(K):
            start = lo - j + K // 2  # index of f(p - j + K//2) for p = 0 in the padded array
            sl = [slice(None)] * 3
            sl[axis] = slice(start, start + n)
            out += k[j] * e[tuple(sl)]
        a = out.astype(np.float32)
    return a


def gaussian_kernel_1d(sigma, normalize=True):
    """mpicbg.imglib.util.Util.createGaussianKernel1DDouble (imglib1 2.0.0-beta, outside the reference
    tree; called at CUDASeparableConvolutionFunctions.java:208)."""
    if sigma <= 0:
        return np.array([0.0, 1.0, 0.0])
    size = max(3, 2 * int(3 * sigma + 0.5) + 1)
    k = np.zeros(size)
    for x in range(size // 2, -1, -1):
        v = np.exp(-(x * x) / (2.0 * sigma * sigma))
        k[size // 2 - x] = v
        k[size // 2 + x] = v
    return k / k.sum() if normalize else k


def cuda_kernels(sigma, supported=(7, 15, 31, 63, 127)):
    """CUDASeparableConvolutionFunctions.getCUDAKernels / getFloatKernelPadded (:199-245)."""
    ks = [gaussian_kernel_1d(s) for s in sigma]
    longest = max(len(k) for k in ks)
    size = min([s for s in supported if longest <= s], default=None)
    if size is None:
        return None
    out = []
    for k in ks:
        p = np.zeros(size, np.float32)
        s0 = (size - len(k)) // 2
        p[s0:s0 + len(k)] = k
        out.append(p)
    return out


# ---- inputs of the loop -------------------------------------------------
def blending_lookup():
    """BlendingRealRandomAccess.java:25-36: lookUp[round(d * 1000)] = (cos((1 - d) pi) + 1) / 2 for the
    accumulated d = 0, 0.001, ... <= 1.0001."""
    lut = np.zeros(1001, np.float64)
    d = 0.0
    while d <= 1.0001:
        i = int(np.floor(d * 1000.0 + 0.5))
        if 0 <= i <= 1000:
            lut[i] = (np.cos((1 - d) * np.pi) + 1) / 2
        d = d + 0.001
    return lut


def blending_weight(t, src_dim_xyz, border, blending, lut=None):
    """BlendingRealRandomAccess.computeWeight (:71-101), float arithmetic as in the Java code.
    t: float32 array [..., 3] of (x, y, z) positions in the view; the interval starts at 0."""
    lut = blending_lookup() if lut is None else lut
    t = np.asarray(t, dtype=F32)
    out = np.ones(t.shape[:-1], F32)
    zero = np.zeros(t.shape[:-1], bool)
    for d in range(3):
        l = t[..., d]
        a = (l - F32(border[d])).astype(F32)
        b = ((F32(src_dim_xyz[d] - 1) - l).astype(F32) - F32(border[d])).astype(F32)
        dist = np.maximum(F32(0), np.minimum(a, b)).astype(F32)
        zero |= dist == 0
        rel = (dist / F32(blending[d])).astype(F32)
        idx = np.floor(rel.astype(np.float64) * 1000.0 + 0.5).astype(np.int64)
        fac = lut[np.clip(idx, 0, 1000)]
        out = np.where(rel < 1, (out.astype(np.float64) * fac).astype(F32), out)
    out[zero] = 0
    return out


def _mirror_single_index(i, n):
    if n == 1:
        return np.zeros_like(i)
    p = 2 * n - 2
    i = np.mod(i, p)
    return np.where(i < n, i, p - i)


def transform_input_and_weights(img, inverse, offset_xyz, out_dim_xyz, border=None, blending=None, min_value=MIN_VALUE):
    """TransformInputAndWeights.loop (TransformInputAndWeights.java:81-114) for a whole bounding box.
    img [z,y,x]; inverse: 3x4 float64 world -> view; returns (transformed [z,y,x], weights or None).
    The n-linear interpolation (imglib2 NLinearInterpolator over Views.extendMirrorSingle, a dependency
    outside the reference tree) is restated in double with one final rounding -- parity unpinned in the
    last bit."""
    inv = np.asarray(inverse, dtype=np.float64).reshape(3, 4)
    X, Y, Z = out_dim_xyz
    sz, sy, sx = img.shape
    zz, yy, xx = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
    s = [(xx + offset_xyz[0]).astype(F32).astype(np.float64), (yy + offset_xyz[1]).astype(F32).astype(np.float64),
         (zz + offset_xyz[2]).astype(F32).astype(np.float64)]
    t = np.stack([(inv[d, 0] * s[0] + inv[d, 1] * s[1] + inv[d, 2] * s[2] + inv[d, 3]).astype(F32) for d in range(3)], axis=-1)
    inside = ((t[..., 0] >= 0) & (t[..., 1] >= 0) & (t[..., 2] >= 0) & (t[..., 0] < F32(sx)) & (t[..., 1] < F32(sy)) & (t[..., 2] < F32(sz)))
    td = t.astype(np.float64)
    f = np.floor(td)
    w = td - f
    i0 = [_mirror_single_index(f[..., d].astype(np.int64), n) for d, n in enumerate((sx, sy, sz))]
    i1 = [_mirror_single_index(f[..., d].astype(np.int64) + 1, n) for d, n in enumerate((sx, sy, sz))]
    a = img.astype(np.float64)

    def at(ix, iy, iz):
        return a[iz, iy, ix]
    c00 = at(i0[0], i0[1], i0[2]) * (1 - w[..., 0]) + at(i1[0], i0[1], i0[2]) * w[..., 0]
    c10 = at(i0[0], i1[1], i0[2]) * (1 - w[..., 0]) + at(i1[0], i1[1], i0[2]) * w[..., 0]
    c01 = at(i0[0], i0[1], i1[2]) * (1 - w[..., 0]) + at(i1[0], i0[1], i1[2]) * w[..., 0]
    c11 = at(i0[0], i1[1], i1[2]) * (1 - w[..., 0]) + at(i1[0], i1[1], i1[2]) * w[..., 0]
    c0 = c00 * (1 - w[..., 1]) + c10 * w[..., 1]
    c1 = c01 * (1 - w[..., 1]) + c11 * w[..., 1]
    val = (c0 * (1 - w[..., 2]) + c1 * w[..., 2]).astype(F32)
    out = np.where(inside, np.maximum(F32(min_value), val), F32(0)).astype(F32)
    weights = None if border is None else blending_weight(t, (sx, sy, sz), border, blending)
    return out, weights


def normalize_weights(weights):
    """WeightNormalizer.ApplyDirectly (WeightNormalizer.java:133-181): w_v = (float)(w_v / sum_v w_v) with a
    double sum -- always applied (the "sumW > 1" test is commented out there), so voxels without any
    weight become NaN; returns (weights, min overlapping views, average overlapping views)."""
    stack = np.stack([np.asarray(w, dtype=F32) for w in weights])
    s = stack.astype(np.float64).sum(axis=0)
    cnt = (stack > 0).sum(axis=0)
    with np.errstate(invalid="ignore", divide="ignore"):
        out = [(w.astype(np.float64) / s).astype(F32) for w in stack]
    return out, int(cnt.min()), float(cnt.mean())


def adjust_for_osem(w, osemspeedup):
    """ProcessForDeconvolution.adjustForOSEM (:351-362): min(1, w * (float)osemspeedup); Math.min keeps NaN."""
    if osemspeedup == 1.0:
        return w
    v = (np.asarray(w, dtype=F32) * F32(osemspeedup)).astype(F32)
    return np.where(np.isnan(v), v, np.minimum(F32(1), v)).astype(F32)


# ---- virtual weights (the dialog's default weight type) ------------------------------------------
def virtual_blending(inverse, offset_xyz, out_dim_xyz, src_dim_xyz, border, blending):
    """The view's blending function seen through TransformedRealRandomAccessibleInterval
    (ProcessForDeconvolution.java:204-208): TransformedInterpolatedRealRandomAccess.applyInverse (:68-95)
    maps the integer world coordinate (voxel + offset, as double) through the inverse registration,
    t0*i00 + t1*i01 + t2*i02 + i03 evaluated left to right in double and stored as float; the float
    position is handed to BlendingRealRandomAccess.computeWeight (:71-101).  Returns float32 [z,y,x]."""
    inv = np.asarray(inverse, dtype=np.float64).reshape(3, 4)
    X, Y, Z = out_dim_xyz
    zz, yy, xx = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
    t0 = (xx + int(offset_xyz[0])).astype(np.float64)
    t1 = (yy + int(offset_xyz[1])).astype(np.float64)
    t2 = (zz + int(offset_xyz[2])).astype(np.float64)
    t = np.stack([(((t0 * inv[d, 0] + t1 * inv[d, 1]) + t2 * inv[d, 2]) + inv[d, 3]).astype(F32) for d in range(3)], axis=-1)
    return blending_weight(t, src_dim_xyz, border, blending)


def compute_sum_image(weights):
    """WeightNormalizer.ComputeSumImage (WeightNormalizer.java:183-232): per voxel the double sum of the
    views' float weights, stored as float where it exceeds 1 and as 1 elsewhere; returns
    (sumWeights, min overlapping views, average overlapping views (global mean))."""
    stack = np.stack([np.asarray(w, dtype=F32) for w in weights])
    s = np.zeros(stack.shape[1:], np.float64)
    for w in stack:  # same order of additions as the cursor loop
        s = s + w.astype(np.float64)
    cnt = (stack > 0).sum(axis=0)
    sum_w = np.where(s > 1, s.astype(F32), F32(1)).astype(F32)
    return sum_w, int(cnt.min()), float(cnt.mean())


def normalizing_weight(w, sum_w, osemspeedup=1.0):
    """NormalizingRandomAccess.get (NormalizingRandomAccess.java:36-45):
    (float) Math.min(1, (double)w / (double)sumW * osemspeedup)."""
    v = np.asarray(w, dtype=F32).astype(np.float64) / np.asarray(sum_w, dtype=F32).astype(np.float64) * float(osemspeedup)
    return np.where(np.isnan(v), v, np.minimum(1.0, v)).astype(F32)


# ---- PSF extraction (ExtractPSF.java) -----------------------------------------
def _sample_linear(img, px, py, pz, mode):
    """n-linear interpolation of img [z,y,x] at double positions (arrays); mode "zero" (Views.extendZero)
    or "periodic" (Views.extendPeriodic).  imglib2's interpolator is outside the reference tree: restated
    in double, one final rounding by the caller (parity unpinned in the last bit)."""
    sz, sy, sx = img.shape
    a = img.astype(np.float64)
    idx, w, ok = [], [], []
    for p, n in ((px, sx), (py, sy), (pz, sz)):
        f = np.floor(p)
        w.append(p - f)
        i0, i1 = f.astype(np.int64), f.astype(np.int64) + 1
        if mode == "periodic":
            idx.append((np.mod(i0, n), np.mod(i1, n)))
            ok.append((np.ones(i0.shape, bool), np.ones(i0.shape, bool)))
        else:
            ok.append(((i0 >= 0) & (i0 < n), (i1 >= 0) & (i1 < n)))
            idx.append((np.clip(i0, 0, n - 1), np.clip(i1, 0, n - 1)))

    def at(bx, by, bz):
        v = a[idx[2][bz], idx[1][by], idx[0][bx]]
        return np.where(ok[0][bx] & ok[1][by] & ok[2][bz], v, 0.0)
    c00 = at(0, 0, 0) * (1 - w[0]) + at(1, 0, 0) * w[0]
    c10 = at(0, 1, 0) * (1 - w[0]) + at(1, 1, 0) * w[0]
    c01 = at(0, 0, 1) * (1 - w[0]) + at(1, 0, 1) * w[0]
    c11 = at(0, 1, 1) * (1 - w[0]) + at(1, 1, 1) * w[0]
    c0 = c00 * (1 - w[1]) + c10 * w[1]
    c1 = c01 * (1 - w[1]) + c11 * w[1]
    return c0 * (1 - w[2]) + c1 * w[2]


def extract_psf_local(img, locations, size_xyz):
    """ExtractPSF.extractPSFLocal (:353-394): float accumulation over the bead locations (x,y,z) in order."""
    X, Y, Z = size_xyz
    zz, yy, xx = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
    psf = np.zeros((Z, Y, X), F32)
    for loc in locations:
        v = _sample_linear(img, xx - X // 2 + float(loc[0]), yy - Y // 2 + float(loc[1]), zz - Z // 2 + float(loc[2]), "periodic")
        psf = (psf + v.astype(F32)).astype(F32)
    return psf


def normalize_psf(psf):
    """ExtractPSF.normalize (:277-295)."""
    a = psf.astype(np.float64)
    return ((a - a.min()) / (a.max() - a.min())).astype(F32)


def transformed_psf_geometry(dim_xyz, model):
    """ExtractPSF.transformPSF (:304-346): (new size, offset) from the bounds of the transformed interval."""
    m = np.asarray(model, dtype=np.float64).reshape(3, 4)
    corners = np.array([[(c >> d & 1) * (dim_xyz[d] - 1.0) for d in range(3)] for c in range(8)])
    t = corners @ m[:, :3].T + m[:, 3]
    size = t.max(axis=0) - t.min(axis=0)
    new = [int(s) + 1 for s in size]
    new = [n + 1 if n % 2 == 0 else n for n in new]
    centre = np.array([float(d // 2) for d in dim_xyz])
    tc = m[:, :3] @ centre + m[:, 3]
    return new, [tc[d] - (new[d] // 2) for d in range(3)]


def transform_psf(psf, model):
    """ExtractPSF.transformPSF + transform (:396-437): resample through the view's model, zero outside."""
    m = np.eye(4)
    m[:3, :4] = np.asarray(model, dtype=np.float64).reshape(3, 4)
    inv = np.linalg.inv(m)
    new, off = transformed_psf_geometry(psf.shape[::-1], model)
    zz, yy, xx = np.meshgrid(np.arange(new[2]), np.arange(new[1]), np.arange(new[0]), indexing="ij")
    t = [xx + off[0], yy + off[1], zz + off[2]]
    p = [inv[d, 0] * t[0] + inv[d, 1] * t[1] + inv[d, 2] * t[2] + inv[d, 3] for d in range(3)]
    return _sample_linear(psf, p[0], p[1], p[2], "zero").astype(F32)


def average_transformed_psf(psfs):
    """ExtractPSF.computeAverageTransformedPSF (:173-215): the PSFs are *added* point-mirrored about their centres
    into an image of the maximal size (loc = psfCentre - loc + avgCentre), zero outside."""
    size = [max(p.shape[2 - d] for p in psfs) for d in range(3)]  # x, y, z
    avg = np.zeros(size[::-1], F32)
    ac = [s // 2 for s in size]
    for p in psfs:
        pz, py, px = p.shape
        pc = (px // 2, py // 2, pz // 2)
        for z in range(pz):
            for y in range(py):
                for x in range(px):
                    lx, ly, lz = pc[0] - x + ac[0], pc[1] - y + ac[1], pc[2] - z + ac[2]
                    if 0 <= lx < size[0] and 0 <= ly < size[1] and 0 <= lz < size[2]:
                        avg[lz, ly, lx] += p[z, y, x]
    return avg


def _conv_zero_same(inp, kernel):
    # new FFTConvolution(Views.extendZero(input), input, Views.extendZero(kernel), kernel, output)
    return convolve_extended(inp, kernel, "zero")


def init_kernels(psfs, psftype, num_threads=8):
    """Returns (kernel1[], kernel2[]) as float32 arrays [z,y,x]."""
    assert psftype in PSFTYPES
    k1 = [np.array(p, dtype=F32, copy=True) for p in psfs]
    k2 = [None] * len(k1)
    V = len(k1)
    for v in range(V):
        k1[v] = norm_img(k1[v], num_threads)  # :165
        if V == 1 or psftype == "INDEPENDENT":  # :174-180
            k2[v] = compute_inverted_kernel(k1[v])
        elif psftype == "EFFICIENT_BAYESIAN":  # :181-244
            tmp = compute_inverted_kernel(k1[v])
            for w in range(V):
                if w == v:
                    continue
                out = _conv_zero_same(compute_inverted_kernel(k1[v]), k1[w])
                out = _conv_zero_same(out, compute_inverted_kernel(k1[w]))
                tmp = (out * tmp).astype(F32)
            k2[v] = norm_img(tmp, num_threads)
        elif psftype == "OPTIMIZATION_I":  # :245-291
            tmp = k1[v].copy()
            for w in range(V):
                if w == v:
                    continue
                out = _conv_zero_same(k1[v], compute_inverted_kernel(k1[w]))
                tmp = (out * tmp).astype(F32)
            tmp = norm_img(tmp, num_threads)
            k2[v] = compute_inverted_kernel(tmp)
        else:  # OPTIMIZATION_II :292-302
            e = compute_exponential_kernel(k1[v], V)
            e = norm_img(e, num_threads)
            k2[v] = compute_inverted_kernel(e)
    return k1, k2


# --------------------------------------------------------------------------------------------
# MVDeconvolution.java:192-235 + FirstIteration.java:63-171
def fuse_first_iteration(imgs):
    """Returns (avg, count_image).  avg = mean over voxels with >= 1 view of the mean of the views
    having data there (double accumulation)."""
    stack = np.stack([np.asarray(i, dtype=F32).astype(np.float64) for i in imgs])
    pos = stack > 0
    cnt = pos.sum(axis=0)
    s = np.where(pos, stack, 0.0).sum(axis=0)
    has = cnt > 0
    total = math.fsum((s[has] / cnt[has]).ravel())
    n = int(has.sum())
    avg = total / n if n > 0 else float("nan")
    return avg, cnt.astype(F32)


# MVDeconvolution.java:473-525
def compute_quotient(psi_blurred, img):
    pb = np.asarray(psi_blurred, dtype=F32)
    im = np.asarray(img, dtype=F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        q = (im / pb).astype(F32)
    return np.where(im > 0, q, F32(1.0)).astype(F32)


# MVDeconvolution.java:705
def tikhonov(value, lam):
    return (np.sqrt(1.0 + 2.0 * lam * value) - 1.0) / lam


# MVDeconvolution.java:582-703 (computeFinalValues + computeNextValue + change)
def compute_final_values(psi, integral, weight, lam, min_value=MIN_VALUE):
    psi = np.asarray(psi, dtype=F32)
    integral = np.asarray(integral, dtype=F32)
    weight = np.broadcast_to(np.asarray(weight, dtype=F32), psi.shape)
    with np.errstate(all="ignore"):
        value = (psi * integral).astype(F32)
        if lam > 0:
            adj_pos = tikhonov(value.astype(np.float64), float(lam)).astype(F32)
        else:
            adj_pos = value
        adjusted = np.where(value > 0, adj_pos, F32(min_value)).astype(F32)
        nxt = np.where(np.isnan(adjusted), F32(min_value), np.maximum(F32(min_value), adjusted)).astype(F32)
        new = (psi + ((nxt - psi).astype(F32) * weight).astype(F32)).astype(F32)
        change = np.abs((new - psi).astype(F32)).astype(F32)
    sum_change = float(np.sum(change.astype(np.float64)))
    max_change = float(np.max(change)) if change.size else -1.0
    return new, sum_change, max_change


def run_deconvolution(imgs, weights, kernel1, kernel2, num_iterations, lam, psi0=None, min_value=MIN_VALUE, conv=None,
                      on_iteration=None):
    """MVDeconvolution constructor (MVDeconvolution.java:73-190) given prepared kernels.
    weights[v]: array or scalar.  Returns (psi, stats[iter][view] = (sum, max), avg).
    on_iteration(i, psi) sees the estimate after every iteration (the reference's debug stack,
    MVDeconvolution.java:139-175), before the final mask."""
    conv = conv or convolve_extended
    V = len(imgs)
    avg, cnt = fuse_first_iteration(imgs)
    if psi0 is None:
        if math.isnan(avg):
            avg = 0.5  # :117-121
        psi = np.full(np.asarray(imgs[0]).shape, F32(avg), dtype=F32)
    else:
        psi = np.array(psi0, dtype=F32, copy=True)
    stats = []
    for _ in range(num_iterations):  # runIteration :333-444
        row = []
        for v in range(V):
            t1 = conv(psi, kernel1[v], "mirror")            # convolve1 :362
            t1 = compute_quotient(t1, imgs[v])              # :371-385
            t2 = conv(t1, kernel2[v], "const", 1.0)         # convolve2 :397
            psi, s, m = compute_final_values(psi, t2, weights[v], lam, min_value)  # :405-436
            row.append((s, m))
        stats.append(row)
        if on_iteration is not None:
            on_iteration(len(stats) - 1, psi)
    psi = np.where(cnt == 0, F32(0), psi).astype(F32)       # :180-187
    return psi, stats, avg


# --------------------------------------------------------------------------------------------
# J/spim/process/cuda/BlockGeneratorFixedSizePrecise.java:25-101; all vectors in (x,y,z) order
def divide_into_blocks(img_size, kernel_size, block_size):
    n = len(img_size)
    eff = [block_size[d] - kernel_size[d] + 1 for d in range(n)]
    if any(e <= 0 for e in eff):
        return None  # :38-42
    loc = [kernel_size[d] // 2 for d in range(n)]
    num = [img_size[d] // eff[d] + (1 if img_size[d] % eff[d] else 0) for d in range(n)]
    blocks = []
    total = 1
    for v in num:
        total *= v
    cur = [0] * n
    for _ in range(total):  # dimension 0 fastest (:68-76)
        eoff = [cur[d] * eff[d] for d in range(n)]
        off = [eoff[d] - kernel_size[d] // 2 for d in range(n)]
        esize = [min(eff[d], img_size[d] - eoff[d]) for d in range(n)]
        blocks.append(dict(blockSize=list(block_size), offset=off, effectiveSize=esize, effectiveOffset=eoff,
                           effectiveLocalOffset=list(loc)))
        for d in range(n):
            cur[d] += 1
            if cur[d] < num[d]:
                break
            cur[d] = 0
    return blocks


# Block.copyBlock (Block.java:114-153, 233-271): block[p] = extended_source(p + offset)
def copy_block(source, block, mode, value=0.0):
    """Block.copyBlock (Block.java:114-153, 233-271): block[x] = extended_source(x + offset) for the B^3 block.
    Reads the block's part of the volume only, like the reference -- the out-of-bounds rule is applied to the faces
    of that part that lie on the volume border (whole-volume padding only for volumes thinner than the halo)."""
    src = np.asarray(source, dtype=F32)
    bs, off = block["blockSize"], block["offset"]
    n = src.shape[::-1]  # (x,y,z)
    lo = [max(0, -off[d]) for d in range(3)]
    hi = [max(0, off[d] + bs[d] - n[d]) for d in range(3)]
    s = [max(0, off[d]) for d in range(3)]
    e = [min(n[d], off[d] + bs[d]) for d in range(3)]
    small = any(e[d] - s[d] <= max(lo[d], hi[d]) or e[d] <= s[d] for d in range(3))
    if small:
        ext = extend(src, lo[::-1], hi[::-1], mode, value)
        t = [off[d] + lo[d] for d in range(3)]
        return np.ascontiguousarray(ext[t[2]:t[2] + bs[2], t[1]:t[1] + bs[1], t[0]:t[0] + bs[0]]).astype(F32)
    sub = src[s[2]:e[2], s[1]:e[1], s[0]:e[0]]  # clipped exactly at the volume faces the block crosses
    return np.ascontiguousarray(extend(sub, lo[::-1], hi[::-1], mode, value)).astype(F32)


# Block.pasteBlock (Block.java:155-195, 313-359)
def paste_block(target, block_data, block):
    eo, es, el = block["effectiveOffset"], block["effectiveSize"], block["effectiveLocalOffset"]
    target[eo[2]:eo[2] + es[2], eo[1]:eo[1] + es[1], eo[0]:eo[0] + es[0]] = \
        block_data[el[2]:el[2] + es[2], el[1]:el[1] + es[1], el[0]:el[0] + es[0]]


def convolve_blockwise(f, kernel, mode, value, block_size, block_conv):
    """The reference's blocked path (MVDeconFFTThreads.java:52-94): copyBlock from the extended
    source, convolve the block with `block_conv(block, kernel)` (any extension), pasteBlock."""
    f = np.asarray(f, dtype=F32)
    ks = list(np.asarray(kernel).shape[::-1])
    blocks = divide_into_blocks(list(f.shape[::-1]), ks, list(block_size))
    out = np.zeros_like(f)
    for b in blocks:
        data = copy_block(f, b, mode, value)
        paste_block(out, block_conv(data, kernel), b)
    return out
