"""Seeded synthetic multi-view data for tests and bench.py.

Small volumes are generated with numpy only (no library call) so that the oracle and the CUDA
path can be fed the same arrays.  Arrays are float32 indexed [z, y, x].
"""
import numpy as np

F32 = np.float32
SEED0 = 14235235  # the only seed constant in the reference (AdjustInput.java:22)


def gaussian_psf(size, sigma_lateral, sigma_axial, angle_deg):
    """Anisotropic Gaussian (long axis = z) rotated about the y axis, sampled on an odd grid.
    size: (kx, ky, kz).  Returns float32 [kz, ky, kx], un-normalised (max 1)."""
    kx, ky, kz = size
    z, y, x = np.meshgrid(np.arange(kz) - kz // 2, np.arange(ky) - ky // 2, np.arange(kx) - kx // 2, indexing="ij")
    a = np.deg2rad(angle_deg)
    xr = np.cos(a) * x + np.sin(a) * z
    zr = -np.sin(a) * x + np.cos(a) * z
    g = np.exp(-0.5 * ((xr / sigma_lateral) ** 2 + (y / sigma_lateral) ** 2 + (zr / sigma_axial) ** 2))
    return g.astype(F32)


def make_psfs(num_views, size, extracted=False, seed=SEED0):
    """PSF_v rotated by v*180/V degrees.  `extracted` adds multiplicative noise U(0.95,1.05) and
    min-max normalisation like ExtractPSF.normalize (ExtractPSF.java:277-295)."""
    k = size[0]
    sl, sa = (1.5, 4.0) if k >= 31 else (1.0, 2.5)
    rng = np.random.default_rng(seed + 1000)
    out = []
    for v in range(num_views):
        p = gaussian_psf(size, sl, sa, v * 180.0 / num_views)
        if extracted:
            p = p * rng.uniform(0.95, 1.05, size=p.shape).astype(F32)
            p = ((p - p.min()) / (p.max() - p.min())).astype(F32)
        out.append(np.ascontiguousarray(p, dtype=F32))
    return out


def make_truth(shape, kind="blob", seed=SEED0):
    """Blob volume (N/4096 Gaussian blobs, sigma U(1.5,4)) or bead phantom (N/32768 beads,
    sigma 0.8) on a 0.01 background.  numpy only; meant for volumes up to ~128^3."""
    rng = np.random.default_rng(seed)
    Z, Y, X = shape
    n = Z * Y * X
    vol = np.full(shape, 0.01, dtype=F32)
    count = max(4, n // (4096 if kind == "blob" else 32768))
    for _ in range(count):
        c = rng.uniform(0, 1, 3) * np.array([Z, Y, X])
        amp = rng.uniform(0.2, 1.0)
        sig = rng.uniform(1.5, 4.0) if kind == "blob" else 0.8
        r = int(np.ceil(3 * sig)) + 1
        z0, z1 = max(0, int(c[0]) - r), min(Z, int(c[0]) + r + 1)
        y0, y1 = max(0, int(c[1]) - r), min(Y, int(c[1]) + r + 1)
        x0, x1 = max(0, int(c[2]) - r), min(X, int(c[2]) + r + 1)
        if z0 >= z1 or y0 >= y1 or x0 >= x1:
            continue
        zz, yy, xx = np.meshgrid(np.arange(z0, z1) - c[0], np.arange(y0, y1) - c[1], np.arange(x0, x1) - c[2], indexing="ij")
        vol[z0:z1, y0:y1, x0:x1] += (amp * np.exp(-0.5 * (zz ** 2 + yy ** 2 + xx ** 2) / sig ** 2)).astype(F32)
    return vol


def degrade(blurred, view, num_views, seed=SEED0, gain=200.0, nodata_fraction=0.12):
    """blurred = conv(truth, PSF_v).  Poisson-like noise (Gaussian approximation, `gain` photons at
    intensity 1), clamp to >= 1e-4, rescale to max 1, and a slab of 12% of x at an end alternating
    by view set to 0 = "no data" (exercises the img > 0 rule, MVDeconvolution.java:496)."""
    rng = np.random.default_rng(seed + 17 * (view + 1))
    b = np.asarray(blurred, dtype=F32)
    noise = rng.standard_normal(b.shape, dtype=F32)
    img = b + np.sqrt(np.maximum(b, 0) / F32(gain)).astype(F32) * noise
    img = np.maximum(img, F32(1e-4))
    img = (img / img.max()).astype(F32)
    X = b.shape[2]
    w = int(round(nodata_fraction * X))
    if w > 0:
        if view % 2 == 0:
            img[:, :, :w] = 0
        else:
            img[:, :, X - w:] = 0
    return np.ascontiguousarray(img, dtype=F32)


def blending_weights(shape, num_views, border=(8, 8, 3), blend_range=12.0):
    """Pre-normalised cosine-blended weights: w_v = blend_v / sum_u blend_u, clamped to <= 1
    (the precomputed-weights branch, WeightNormalizer.java:117-164; blend after
    BlendingRealRandomAccess.java:71-101 with the view box shifted per view along x)."""
    Z, Y, X = shape
    ws = []
    for v in range(num_views):
        w = np.ones(shape, dtype=np.float64)
        for ax, n, b in ((0, Z, border[2]), (1, Y, border[1]), (2, X, border[0])):
            pos = np.arange(n, dtype=np.float64)
            shift = (v - (num_views - 1) / 2.0) * 0.05 * n if ax == 2 else 0.0
            d = np.minimum(pos - shift, (n - 1) - (pos - shift)) - b
            rel = np.clip(d / blend_range, 0.0, 1.0)
            ramp = 0.5 * (np.cos((1.0 - rel) * np.pi) + 1.0)
            sh = [1, 1, 1]
            sh[ax] = n
            w = w * ramp.reshape(sh)
        ws.append(w)
    total = np.maximum(np.sum(ws, axis=0), 1e-12)
    return [np.ascontiguousarray(np.minimum(1.0, w / total), dtype=F32) for w in ws]


def make_dataset(shape, num_views, psf_size, kind="blob", extracted=False, seed=SEED0, convolve=None):
    """Returns dict(truth, psfs, imgs).  `convolve(f, k)` must implement conv with mirror-single
    extension; the caller supplies it (oracle in tests, library in bench)."""
    truth = make_truth(shape, kind, seed)
    psfs = make_psfs(num_views, psf_size, extracted, seed)
    imgs = []
    for v in range(num_views):
        k = (psfs[v].astype(np.float64) / psfs[v].astype(np.float64).sum()).astype(F32)
        imgs.append(degrade(convolve(truth, k), v, num_views, seed))
    return dict(truth=truth, psfs=psfs, imgs=imgs)
