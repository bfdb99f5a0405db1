"""CPU tests of the oracle (oracle/oracle_np.py) -- no GPU needed.

The reference ships no tests or golden data for this path, so the oracle
is pinned against (a) the known answers derivable from the reference's own `main` smoke programs
(Block.java:367-402, MVDeconvolution.java:707-714), (b) an FFT-free brute-force restatement of
the convolution formula, (c) the invariants the reference's structure implies (tiling invariance,
fixed point, delta placement), and (d) the frozen golden fixture tests/golden/loop_3view_32.npz.
"""
import math
import os

import numpy as np
import pytest

import oracle_np as o
from conftest import rel_l2

F32 = np.float32


# ---- (a) known answers from the reference's main() programs -----------------------------------
def test_block_kat_from_block_main():
    """Block.java:367-402: image 1024x1024, block 384x384, kernel (16, 32)."""
    blocks = o.divide_into_blocks([1024, 1024], [16, 32], [384, 384])
    assert len(blocks) == 9  # 3 x 3
    eff = [384 - 16 + 1, 384 - 32 + 1]
    assert eff == [369, 353]
    b0, b1, b3, last = blocks[0], blocks[1], blocks[3], blocks[-1]
    assert b0["effectiveOffset"] == [0, 0] and b0["offset"] == [-8, -16] and b0["effectiveSize"] == eff
    assert b1["effectiveOffset"] == [369, 0]  # dimension 0 fastest
    assert b3["effectiveOffset"] == [0, 353]
    assert last["effectiveOffset"] == [738, 706] and last["effectiveSize"] == [286, 318]
    assert all(b["blockSize"] == [384, 384] and b["effectiveLocalOffset"] == [8, 16] for b in blocks)
    # the effective regions partition the image exactly
    cover = np.zeros((1024, 1024), np.int32)
    for b in blocks:
        eo, es = b["effectiveOffset"], b["effectiveSize"]
        cover[eo[1]:eo[1] + es[1], eo[0]:eo[0] + es[0]] += 1
    assert (cover == 1).all()


def test_block_counts_at_baseline_configs():
    """cfg3 B=256,K=31 -> E=226, 5x5x3 = 75 blocks, last effective (120,120,60)."""
    blocks = o.divide_into_blocks([1024, 1024, 512], [31, 31, 31], [256, 256, 256])
    assert len(blocks) == 75
    assert blocks[-1]["effectiveSize"] == [120, 120, 60]
    assert len(o.divide_into_blocks([512, 512, 256], [31] * 3, [256] * 3)) == 18
    assert len(o.divide_into_blocks([2048, 2048, 1024], [31] * 3, [256] * 3)) == 500
    assert len(o.divide_into_blocks([2048, 2048, 1024], [31] * 3, [512] * 3)) == 75
    assert o.divide_into_blocks([64, 64, 64], [31] * 3, [30] * 3) is None  # :38-42 returns null


def test_tikhonov_table():
    """MVDeconvolution.java:705-714: tikhonov(d, 0.0006) for d in [0, 10) and 10000*d."""
    lam = 0.0006
    d = 0.0
    while d < 10:
        for v in (d * 10000, d):
            want = (math.sqrt(1.0 + 2.0 * lam * v) - 1.0) / lam
            assert o.tikhonov(np.float64(v), lam) == pytest.approx(want, rel=1e-15, abs=0)
        d += 0.1
    # the regulariser shrinks, is monotone and tends to the identity for small values
    v = np.linspace(1e-3, 50, 1000)
    t = o.tikhonov(v, 0.006)
    assert (t < v).all() and (np.diff(t) > 0).all()
    assert o.tikhonov(np.float64(1e-6), 0.006) == pytest.approx(1e-6, rel=1e-5)


def test_update_rule_scalar_cases():
    """computeNextValue (MVDeconvolution.java:671-703) on hand-evaluated cases."""
    psi = np.array([0.5, 0.5, 0.5, 0.5, 2.0], F32)
    integral = np.array([2.0, -1.0, np.nan, 0.0, 1e-6], F32)
    w = np.array([1.0, 1.0, 1.0, 0.25, 0.5], F32)
    new, s, m = o.compute_final_values(psi, integral, w, 0.0)
    # value>0 -> value ; value<=0 -> minValue ; NaN -> minValue ; weights blend
    nxt = np.array([1.0, 1e-4, 1e-4, 1e-4, 1e-4], F32)
    want = (psi + ((nxt - psi).astype(F32) * w).astype(F32)).astype(F32)  # float arithmetic, as in Java (:702)
    assert want[0] == 1.0 and abs(float(want[1]) - 1e-4) < 1e-7 and want[3] == F32(0.5) + (F32(1e-4) - F32(0.5)) * F32(0.25)
    np.testing.assert_array_equal(new, want)
    assert m == pytest.approx(float(np.abs(want - psi).max()))
    assert s == pytest.approx(float(np.abs((want - psi).astype(F32)).astype(np.float64).sum()))
    new, _, _ = o.compute_final_values(np.array([0.5], F32), np.array([2.0], F32), F32(1.0), 0.006)
    assert new[0] == F32((math.sqrt(1.0 + 2.0 * 0.006 * 1.0) - 1.0) / 0.006)


def test_quotient_rule():
    """MVDeconvolution.java:496-499: img > 0 ? img / blurred : 1; no guard on blurred == 0."""
    q = o.compute_quotient(np.array([2.0, 0.0, 4.0, 1.0], F32), np.array([1.0, 1.0, 0.0, -1.0], F32))
    assert q[0] == F32(0.5) and np.isinf(q[1]) and q[2] == 1.0 and q[3] == 1.0


# ---- kernel preparation quirks -----------------------------------------------------------------
def test_sum_img_double_counts_first_portion():
    """AdjustInput.java:62,72,93-97."""
    img = np.arange(1, 101, dtype=F32)
    T = 4
    chunk = 100 // (2 * T)
    assert o.sum_img(img, T) == pytest.approx(float(img.sum()) + float(img[:chunk].sum()))
    k = o.norm_img(np.ones((5, 5, 5), F32), 8)
    assert float(k.astype(np.float64).sum()) < 1.0  # normalised kernels sum to slightly < 1


def test_mirror_odd_is_flip_even_keeps_middle_pair():
    a = np.arange(7, dtype=F32).reshape(1, 1, 7)
    np.testing.assert_array_equal(o.mirror(a, 0)[0, 0], a[0, 0, ::-1])
    b = np.arange(6, dtype=F32).reshape(1, 1, 6)
    # Mirror.java:72: sizeMirrorH = dim/2 and the loop runs p <= sizeMirrorH... the middle pair is swapped twice
    np.testing.assert_array_equal(o.mirror(b, 0)[0, 0], np.array([5, 4, 2, 3, 1, 0], F32))


@pytest.mark.parametrize("psftype", o.PSFTYPES)
def test_kernel2_properties(psftype):
    from spim_registration_b200 import synthetic
    psfs = synthetic.make_psfs(3, (7, 5, 9), extracted=True)
    k1, k2 = o.init_kernels(psfs, psftype, 8)
    for v in range(3):
        assert k1[v].dtype == F32 and k2[v].dtype == F32 and k2[v].shape == k1[v].shape
        assert 0.9 < float(k1[v].astype(np.float64).sum()) < 1.0
    if psftype == "INDEPENDENT":
        for v in range(3):
            np.testing.assert_array_equal(k2[v], k1[v][::-1, ::-1, ::-1])
    if psftype == "OPTIMIZATION_II":
        e = (k1[0] * k1[0]).astype(F32)
        e = (e * k1[0]).astype(F32)
        np.testing.assert_array_equal(k2[0], o.norm_img(e, 8)[::-1, ::-1, ::-1])
    if psftype in ("EFFICIENT_BAYESIAN", "OPTIMIZATION_I"):
        assert 0.9 < float(k2[0].astype(np.float64).sum()) < 1.0


def test_single_view_uses_flipped_kernel_for_every_type():
    """MVDeconFFT.java:174-180: numViews == 1 -> kernel2 = flip(kernel1) whatever the type."""
    from spim_registration_b200 import synthetic
    psf = synthetic.make_psfs(1, (5, 5, 5))
    for t in o.PSFTYPES:
        k1, k2 = o.init_kernels(psf, t, 8)
        np.testing.assert_array_equal(k2[0], k1[0][::-1, ::-1, ::-1])


# ---- (b) convolution formula against brute force ----------------------------------------------
@pytest.mark.parametrize("mode", ["zero", "const", "mirror"])
@pytest.mark.parametrize("shape,ks", [((6, 7, 8), (3, 5, 3)), ((5, 4, 9), (4, 2, 6)), ((3, 3, 3), (7, 7, 7))])
def test_fft_oracle_equals_direct_sum(mode, shape, ks):
    rng = np.random.default_rng(5)
    f = rng.random(shape, dtype=F32)
    k = rng.random(ks, dtype=F32)
    a = o.convolve_extended(f, k, mode, 1.0)
    b = o.convolve_direct(f, k, mode, 1.0)
    np.testing.assert_allclose(a, b, rtol=2e-6, atol=1e-6)


def test_delta_places_kernel_unflipped():
    """Convolution, not correlation (SURVEY.md 8c caveat): delta at p -> kernel[j] at p + j - K//2."""
    k = np.arange(1, 3 * 5 * 7 + 1, dtype=F32).reshape(3, 5, 7)
    d = np.zeros((9, 11, 13), F32)
    d[4, 5, 6] = 1
    out = o.convolve_extended(d, k, "zero")
    np.testing.assert_allclose(out[3:6, 3:8, 3:10], k, atol=1e-4)


def test_mirror_single_extension_indices():
    a = np.arange(5, dtype=F32).reshape(1, 1, 5)
    e = o.extend(a, (0, 0, 3), (0, 0, 3), "mirror")
    np.testing.assert_array_equal(e[0, 0], np.array([3, 2, 1, 0, 1, 2, 3, 4, 3, 2, 1], F32))
    e = o.extend(a, (0, 0, 9), (0, 0, 9), "mirror")  # halo wider than the image: period 2n-2
    idx = np.arange(-9, 14)
    p = np.mod(idx, 8)
    np.testing.assert_array_equal(e[0, 0], np.where(p >= 5, 8 - p, p).astype(F32))


# ---- (c) structural invariants ---------------------------------------------------------------
@pytest.mark.parametrize("mode,val", [("mirror", 0.0), ("const", 1.0)])
def test_tiling_invariance(mode, val):
    """SURVEY.md 8a invariant (i): reference-style blocks with ANY in-block extension == whole volume."""
    rng = np.random.default_rng(9)
    f = rng.random((20, 26, 30), dtype=F32)
    k = rng.random((5, 7, 3), dtype=F32)
    whole = o.convolve_extended(f, k, mode, val)
    tiled = o.convolve_blockwise(f, k, mode, val, [16, 14, 12], lambda blk, kk: o.convolve_extended(blk, kk, "zero"))
    assert rel_l2(tiled, whole) < 1e-6


def test_copy_paste_block_roundtrip_is_identity_on_effective_regions():
    rng = np.random.default_rng(2)
    f = rng.random((10, 12, 14), dtype=F32)
    blocks = o.divide_into_blocks([14, 12, 10], [3, 5, 3], [8, 8, 8])
    out = np.full_like(f, -1)
    for b in blocks:
        o.paste_block(out, o.copy_block(f, b, "mirror"), b)
    np.testing.assert_array_equal(out, f)  # bit exact: pure index arithmetic


def test_fixed_point_of_independent_update():
    from spim_registration_b200 import synthetic
    truth = synthetic.make_truth((14, 16, 18), "blob", 5) + F32(0.05)
    k = np.random.default_rng(1).random((5, 5, 5), dtype=F32)
    k = (k / k.sum()).astype(F32)
    img = o.convolve_extended(truth, k, "mirror")
    psi, _, _ = o.run_deconvolution([img], [F32(1.0)], [k], [k[::-1, ::-1, ::-1].copy()], 3, 0.0, psi0=truth)
    assert rel_l2(psi, truth) < 2e-5


def test_first_iteration_average_and_mask():
    a = np.array([[[0.0, 2.0, 4.0, 0.0]]], F32)
    b = np.array([[[0.0, 4.0, 0.0, 0.0]]], F32)
    avg, cnt = o.fuse_first_iteration([a, b])
    assert avg == pytest.approx((3.0 + 4.0) / 2) and cnt.ravel().tolist() == [0, 2, 1, 0]
    k = np.ones((1, 1, 1), F32)
    psi, stats, _ = o.run_deconvolution([a, b], [F32(0.5), F32(0.5)], [k, k], [k, k], 1, 0.0)
    assert psi[0, 0, 0] == 0 and psi[0, 0, 3] == 0 and (psi[0, 0, 1:3] > 0).all()
    avg, _ = o.fuse_first_iteration([np.zeros((2, 2, 2), F32)])
    assert math.isnan(avg)  # -> 0.5 in run_deconvolution (MVDeconvolution.java:117-121)


# ---- (d) the golden fixture is reproducible ---------------------------------------------------
def test_golden_fixture_reproduced_by_oracle():
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "loop_3view_32.npz"))
    V = int(g["num_views"])
    psfs = [g["psf%d" % v] for v in range(V)]
    k1, k2 = o.init_kernels(psfs, str(g["psftype"]), 8)
    for v in range(V):
        np.testing.assert_array_equal(k1[v], g["k1_%d" % v])
        np.testing.assert_allclose(k2[v], g["k2_%d" % v], rtol=1e-6, atol=1e-12)
    psi, stats, avg = o.run_deconvolution([g["img%d" % v] for v in range(V)], [g["weight%d" % v] for v in range(V)],
                                          [g["k1_%d" % v] for v in range(V)], [g["k2_%d" % v] for v in range(V)],
                                          int(g["iterations"]), float(g["lam"]))
    assert avg == pytest.approx(float(g["avg"]), rel=1e-12)
    assert rel_l2(psi, g["psi"]) < 1e-6


# ---- the separable-convolution hook (CUDASeparableConvolution.java:13-17) -------------------------
@pytest.mark.parametrize("mode,nd_mode", [("zero", "constant"), ("clamp", "nearest"), ("const", "constant")])
def test_separable_convolve_against_scipy(mode, nd_mode):
    """Per-axis in-place passes, centre K // 2 also for even kernels, each pass with its own extension."""
    from scipy import ndimage
    rng = np.random.default_rng(3)
    f = rng.random((9, 11, 13), dtype=np.float32)
    ks = {2: rng.random(7), 1: rng.random(4), 0: rng.random(5)}
    got = o.separable_convolve(f, ks[2], ks[1], ks[0], mode, 0.75)
    ref = f
    for ax in (2, 1, 0):
        ref = ndimage.convolve1d(ref.astype(np.float64), ks[ax], axis=ax, mode=nd_mode, cval=0.75 if mode == "const" else 0.0).astype(np.float32)
    np.testing.assert_allclose(got, ref, rtol=0, atol=2e-6)
    # skipping an axis == convolving it with a delta
    a = o.separable_convolve(f, ks[2], None, ks[0], mode, 0.75)
    b = o.separable_convolve(f, ks[2], np.array([0.0, 1.0, 0.0]), ks[0], mode, 0.75)
    np.testing.assert_array_equal(a, b)


def test_separable_equals_full_3d_convolution_for_rank1_kernels():
    """A rank-1 kernel through the three 1-D passes == the 3-D convolution of the loop's oracle (zero extension)."""
    rng = np.random.default_rng(4)
    f = rng.random((10, 12, 14), dtype=np.float32)
    kx, ky, kz = rng.random(5), rng.random(3), rng.random(7)
    k3 = (kz[:, None, None] * ky[None, :, None] * kx[None, None, :]).astype(np.float32)
    a = o.separable_convolve(f, kx, ky, kz, "zero")
    b = o.convolve_extended(f, k3, "zero")
    assert rel_l2(a, b) <= 2e-6


def test_gaussian_kernels_of_the_hook():
    """Util.createGaussianKernel1DDouble + getCUDAKernels: size rule, symmetry, normalisation, centred padding."""
    k = o.gaussian_kernel_1d(1.8)
    assert len(k) == 2 * int(3 * 1.8 + 0.5) + 1 == 11 and abs(k.sum() - 1) < 1e-12 and np.allclose(k, k[::-1])
    assert list(o.gaussian_kernel_1d(0.0)) == [0.0, 1.0, 0.0] and len(o.gaussian_kernel_1d(0.2)) == 3
    ks = o.cuda_kernels([1.8, 1.8, 2.5])
    assert [len(x) for x in ks] == [31, 31, 31]  # 17 taps for sigma 2.5 -> next supported size
    assert ks[0][:10].sum() == 0 and ks[0][21:].sum() == 0 and abs(ks[0].sum() - 1) < 1e-6
    assert o.cuda_kernels([30.0]) is None or len(o.cuda_kernels([30.0])[0]) == 127
    assert o.cuda_kernels([50.0]) is None  # 301 taps: "Kernel bigger than maximally supported size"
    import spim_registration_b200 as m
    for a, b in zip(m.CUDASeparableConvolutionFunctions.getCUDAKernels([1.8, 1.8, 2.5]), ks):
        np.testing.assert_array_equal(a, b)


# ---- inputs of the loop: resampling, blending weights, weight normalisation ------
def test_blending_weight_known_values():
    """BlendingRealRandomAccess.computeWeight: 0 on and outside the border, cosine ramp of `blending`
    pixels, 1 in the interior, product over the dimensions; the lookup table quantises the ramp to 1/1000."""
    lut = o.blending_lookup()
    assert lut[0] == 0 and abs(lut[500] - 0.5) < 1e-12 and lut[1000] == 1 and np.all(np.diff(lut) >= 0)
    dim, border, blend = (101, 81, 61), (5.0, 0.0, 0.0), (10.0, 20.0, 12.0)
    t = np.array([[50, 40, 30], [5, 40, 30], [4, 40, 30], [10, 40, 30], [50, 0, 30], [50, 10, 30], [95, 40, 30], [50, 40, 60], [7.5, 5, 30]], np.float32)
    w = o.blending_weight(t, dim, border, blend)
    ramp = lambda d: (np.cos((1 - d) * np.pi) + 1) / 2
    assert w[0] == 1 and w[1] == 0 and w[2] == 0 and w[4] == 0 and w[7] == 0
    assert abs(w[3] - ramp(0.5)) < 2e-3 and abs(w[5] - ramp(0.5)) < 2e-3 and abs(w[6] - ramp(0.0)) < 1e-6  # x = 95 = 100 - 5: on the border
    assert abs(w[8] - ramp(0.25) * ramp(0.25)) < 4e-3


def test_transform_identity_translation_and_outside():
    rng = np.random.default_rng(5)
    img = rng.random((6, 7, 8), dtype=np.float32) + np.float32(0.01)
    ident = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0]], float)
    t, w = o.transform_input_and_weights(img, ident, (0, 0, 0), (8, 7, 6))
    np.testing.assert_array_equal(t, img)           # integer positions: interpolation reproduces the samples
    assert w is None
    # the bounding box starts at (-2, 1, 0): voxel x of the box is view x - 2 ... and half-voxel shifts average neighbours
    t2, _ = o.transform_input_and_weights(img, ident, (-2, 1, 0), (8, 7, 6))
    np.testing.assert_array_equal(t2[:, :-1, 2:], img[:, 1:, :-2])
    assert (t2[:, :, :2] == 0).all() and (t2[:, -1, :] == 0).all()   # outside the view: 0 = no data
    half = ident.copy()
    half[0, 3] = 0.5
    t3, _ = o.transform_input_and_weights(img, half, (0, 0, 0), (8, 7, 6))
    np.testing.assert_allclose(t3[:, :, :-1], (img[:, :, :-1].astype(np.float64) + img[:, :, 1:]) / 2, rtol=1e-6)
    np.testing.assert_allclose(t3[:, :, -1], (img[:, :, -1].astype(np.float64) + img[:, :, -2]) / 2, rtol=1e-6)  # mirror-single beyond the last sample
    dark = np.zeros((3, 3, 3), np.float32)
    t4, _ = o.transform_input_and_weights(dark, ident, (0, 0, 0), (3, 3, 3))
    assert (t4 == o.MIN_VALUE).all()                # inside the view the value never drops to the "no data" marker


def test_weight_normalisation_and_osem():
    w = [np.array([0.5, 0.2, 0.0, 0.0], np.float32), np.array([1.0, 0.2, 0.3, 0.0], np.float32)]
    n, mn, avg = o.normalize_weights(w)
    np.testing.assert_allclose(n[0][:3], [1 / 3, 0.5, 0.0], rtol=1e-6)
    np.testing.assert_allclose(n[1][:3], [2 / 3, 0.5, 1.0], rtol=1e-6)
    assert np.isnan(n[0][3]) and np.isnan(n[1][3])  # 0 / 0, exactly what WeightNormalizer.java:171 does
    assert mn == 0 and abs(avg - 5 / 4) < 1e-12
    a = o.adjust_for_osem(n[1], 2.0)
    np.testing.assert_allclose(a[:3], [1.0, 1.0, 1.0])
    assert np.isnan(a[3]) and o.adjust_for_osem(n[1], 1.0) is n[1]


# ---- PSF extraction (ExtractPSF.java) -------------------------------------------
def test_virtual_weights_known_answers():
    """WeightType.VIRTUAL_WEIGHTS: identity registration -> the plain blending image; sum image = max(1, sum);
    NormalizingRandomAccess = min(1, w / sumW * osem); with two identical views each gets 1/2 in the plateau."""
    dims = (40, 30, 20)
    ident = np.eye(4)[:3, :4]
    b = o.virtual_blending(ident, (0, 0, 0), dims, dims, (0, 0, 0), (6, 6, 6))
    assert b.shape == dims[::-1] and b.dtype == np.float32
    assert b[10, 15, 20] == 1.0 and b[0, 15, 20] == 0.0 and b[10, 15, 39] == 0.0
    # position convention: shifted bounding box and view
    shifted = o.virtual_blending(np.array([[1, 0, 0, -5.0], [0, 1, 0, 0], [0, 0, 1, 0]]), (5, 0, 0), dims, dims, (0, 0, 0), (6, 6, 6))
    np.testing.assert_array_equal(shifted, b)
    s, mn, avg = o.compute_sum_image([b, b])
    assert s.min() == 1.0 and s[10, 15, 20] == 2.0 and mn == 0 and 0 < avg < 2
    w = o.normalizing_weight(b, s)
    assert w[10, 15, 20] == 0.5 and w[0, 15, 20] == 0.0
    # in the ramp the sum stays <= 1 until both weights exceed 1/2: there the weight is not normalised
    ramp = (s == 1) & (b > 0)
    np.testing.assert_array_equal(w[ramp], b[ramp])
    np.testing.assert_array_equal(o.normalizing_weight(b, s, 3.0), np.minimum(np.float32(1), (b.astype(np.float64) / s * 3.0)).astype(np.float32))
    # same positions as the TransformInputAndWeights route
    _, w2 = o.transform_input_and_weights(np.ones(dims[::-1], np.float32), ident, (0, 0, 0), dims, (0, 0, 0), (6, 6, 6))
    np.testing.assert_array_equal(w2, b)


def test_extract_psf_known_answers():
    rng = np.random.default_rng(31)
    img = rng.random((12, 14, 16), dtype=np.float32)
    # one bead at an integer position: the PSF is the window around it (periodic wrap at the border)
    psf = o.extract_psf_local(img, [(8, 7, 6)], (5, 5, 3))
    np.testing.assert_array_equal(psf, img[5:8, 5:10, 6:11])
    wrap = o.extract_psf_local(img, [(0, 0, 0)], (3, 3, 3))
    np.testing.assert_array_equal(wrap, np.roll(img, (1, 1, 1), (0, 1, 2))[:3, :3, :3])
    # two beads add; half-voxel positions average neighbours
    two = o.extract_psf_local(img, [(8, 7, 6), (4, 5, 6)], (5, 5, 3))
    np.testing.assert_allclose(two, img[5:8, 5:10, 6:11] + img[5:8, 3:8, 2:7], rtol=1e-6)
    half = o.extract_psf_local(img, [(8.5, 7, 6)], (5, 5, 3))
    np.testing.assert_allclose(half, (img[5:8, 5:10, 6:11].astype(np.float64) + img[5:8, 5:10, 7:12]) / 2, rtol=1e-6)
    n = o.normalize_psf(psf)
    assert n.min() == 0 and n.max() == 1


def test_transform_psf_geometry_and_identity():
    psf = np.random.default_rng(32).random((7, 9, 11), dtype=np.float32)
    ident = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0]], float)
    size, off = o.transformed_psf_geometry((11, 9, 7), ident)
    assert size == [11, 9, 7] and off == [0.0, 0.0, 0.0]
    np.testing.assert_array_equal(o.transform_psf(psf, ident), psf)
    # z scaling 2.5 (anisotropic stack): z extent (7-1)*2.5 = 15 -> 16 -> odd 17; the centre voxel stays the centre
    scale = np.diag([1.0, 1.0, 2.5, 1.0])[:3]
    size, off = o.transformed_psf_geometry((11, 9, 7), scale)
    assert size == [11, 9, 17] and off[2] == 3 * 2.5 - 8
    t = o.transform_psf(psf, scale)
    assert t.shape == (17, 9, 11)
    np.testing.assert_allclose(t[8], psf[3], rtol=1e-6)  # centre plane is reproduced
    even = np.random.default_rng(33).random((6, 6, 6), dtype=np.float32)
    assert o.transform_psf(even, ident).shape == (7, 7, 7)  # even sizes become odd ("works if the kernel is even")


def test_average_transformed_psf_mirrors_about_the_centre():
    a = np.zeros((3, 3, 5), np.float32)
    a[1, 1, 4] = 1.0                                   # one voxel right of the centre (x = 2) by 2
    b = np.zeros((5, 3, 3), np.float32)
    b[0, 1, 1] = 2.0
    avg = o.average_transformed_psf([a, b])
    assert avg.shape == (5, 3, 5)
    assert avg[2, 1, 0] == 1.0 and avg[4, 1, 2] == 2.0 and avg.sum() == 3.0   # point-mirrored: centre - loc + avgCentre
    import spim_registration_b200 as m
    e = m.ExtractPSF()
    e.pointSpreadFunctions = [a, b]
    np.testing.assert_array_equal(e.computeAverageTransformedPSF(), avg)


def test_cpu_baseline_port_matches_oracle():
    """The timed CPU port (float32 pocketfft, 64^3 blocks, threaded elementwise passes) computes the oracle's loop."""
    import cpu_baseline as cb
    from spim_registration_b200 import synthetic
    shape, V = (40, 72, 90), 2
    ds = synthetic.make_dataset(shape, V, (7, 7, 7), convolve=lambda f, k: o.convolve_extended(f, k, "mirror"))
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    rng = np.random.default_rng(4)
    w = [(rng.random(shape, dtype=np.float32) * np.float32(0.5) + np.float32(0.2)) for _ in range(V)]
    psi0 = np.full(shape, 0.3, np.float32)
    ref, _, _ = o.run_deconvolution(ds["imgs"], w, k1, k2, 2, 0.006, psi0=psi0.copy())
    got, seconds = cb.run_iterations(ds["imgs"], w, k1, k2, 2, 0.006, [64, 64, 64], psi0.copy(), workers=3)
    assert seconds > 0
    err = np.linalg.norm(got.astype(np.float64) - ref) / np.linalg.norm(ref.astype(np.float64))
    assert err <= 1e-4


# ---- the second, independent restatement (oracle/cpu_ref.cpp): the reference's own FFT route in float32 --------------
def _cpu_ref():
    import cpu_ref
    cpu_ref.load()
    return cpu_ref


def test_prime_factor_sizes_of_the_fft_padding():
    """Valid Mines-JTK lengths are products of mutually prime factors from {2,3,4,5,7,8,9,11,13,16}
    (FFTMethods.java:1279-1291 pads every dimension to one of them; x to an even one)."""
    lib = _cpu_ref().load()
    assert [lib.ref_nfft_small_complex(n) for n in (1, 17, 32, 33, 142, 286, 287)] == [1, 18, 33, 33, 143, 286, 308]
    assert lib.ref_nfft_small_complex(32) == 33          # 32 = 2^5 is not valid: the largest power of two is 16
    assert [lib.ref_nfft_small_real(n) for n in (142, 286, 45)] == [144, 286, 48]


@pytest.mark.parametrize("shape,ks", [((12, 14, 17), (3, 5, 7)), ((20, 9, 11), (4, 6, 5)), ((6, 7, 8), (9, 9, 9))])
@pytest.mark.parametrize("mode", ["zero", "const", "mirror"])
def test_fft_route_without_conjugate_is_the_convolution_formula(shape, ks, mode):
    """Restating FFTConvolution.convolve (pad, rotate the kernel to its centre K/2, multiply, crop) WITHOUT the complex
    conjugate reproduces conv(ext(f), k)(x) = sum_j ext(f)(x + K/2 - j) k(j) -- the formula oracle_np.py evaluates in
    float64 and the brute-force sum -- for all three out-of-bounds rules, odd and even kernels."""
    cr = _cpu_ref()
    rng = np.random.default_rng(5)
    f = rng.random(shape, dtype=np.float32) + np.float32(0.1)
    k = rng.random(ks, dtype=np.float32)
    got, pad = cr.fft_convolve(f, k, mode, 0.7, conjugate=False)
    assert all(p >= n + q - 1 for p, n, q in zip(pad, shape[::-1], ks[::-1]))
    assert rel_l2(got, o.convolve_extended(f, k, mode, 0.7)) <= 2e-6
    assert rel_l2(got, o.convolve_direct(f, k, mode, 0.7)) <= 2e-6


def test_conjugate_switch_is_correlation_and_which_setting_the_path_needs():
    """The vendored copy defaults complexConjugate = true (J/fftMethods/FFTConvolution.java:95, :528-531).  With the
    kernel spectrum conjugated the same route computes the CORRELATION sum_j ext(f)(x - K/2 + j) k(j), i.e. the
    convolution with the kernel mirrored about its centre.  Which setting the deconvolution path needs follows from
    the reference's own construction: PSFs are averaged bead IMAGES (ExtractPSF.java:353-394), so an observed view is
    conv(truth, psf) in image orientation; conv1 must reproduce it and conv2 uses the explicitly mirrored kernel
    (computeInvertedKernel, MVDeconFFT.java:315-323).  Only without the conjugate is the truth a fixed point of the
    iteration for an asymmetric PSF; with it the estimate drifts away."""
    cr = _cpu_ref()
    rng = np.random.default_rng(6)
    f = rng.random((10, 12, 14), dtype=np.float32)
    k = rng.random((5, 3, 7), dtype=np.float32)
    corr, _ = cr.fft_convolve(f, k, "zero", conjugate=True)
    flipped = np.ascontiguousarray(k[::-1, ::-1, ::-1])
    assert rel_l2(corr, o.convolve_extended(f, flipped, "zero")) <= 2e-6
    assert rel_l2(corr, o.convolve_extended(f, k, "zero")) > 0.05
    # fixed point: img = physical blur of the truth by an asymmetric PSF (a one-sided tail along x)
    shape = (18, 20, 22)
    truth = (rng.random(shape, dtype=np.float32) + np.float32(0.5))
    psf = np.zeros((5, 5, 5), np.float32)
    psf[2, 2, 2:] = (0.5, 0.3, 0.2)
    img = o.convolve_direct(truth, psf, "mirror")
    k1 = [psf]
    k2 = [o.compute_inverted_kernel(psf)]
    drift = {}
    for conj in (False, True):
        psi, _ = cr.run_deconvolution([img], [np.float32(1.0)], k1, k2, 3, 0.0, truth, conjugate=conj)
        inner = (slice(4, -4),) * 3   # away from the border, where ext(q) = 1 differs from the interior anyway
        drift[conj] = rel_l2(psi[inner], truth[inner])
    assert drift[False] <= 5e-6
    assert drift[True] > 1e-2


def test_cpp_loop_equals_numpy_oracle():
    """Two independent restatements of the loop (float32 prime-factor-padded FFT route in C++ vs float64 formula in
    numpy) agree to float rounding over several iterations, statistics included."""
    cr = _cpu_ref()
    from spim_registration_b200 import synthetic
    shape, V, iters = (20, 24, 28), 3, 4
    ds = synthetic.make_dataset(shape, V, (7, 7, 7), convolve=lambda a, b: o.convolve_extended(a, b, "mirror"))
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    weights = synthetic.blending_weights(shape, V)
    ref, stats, avg = o.run_deconvolution(ds["imgs"], weights, k1, k2, iters, 0.006)
    psi0 = np.full(shape, np.float32(avg), np.float32)
    got, gstats = cr.run_deconvolution(ds["imgs"], weights, k1, k2, iters, 0.006, psi0)
    cnt = sum((np.asarray(i) > 0) for i in ds["imgs"])
    got = np.where(cnt == 0, np.float32(0), got)
    assert rel_l2(got, ref) <= 2e-6
    np.testing.assert_allclose(np.array(gstats)[:, :, 0], np.array([[s for s, _ in r] for r in stats]), rtol=2e-5)
    np.testing.assert_allclose(np.array(gstats)[:, :, 1], np.array([[m for _, m in r] for r in stats]), rtol=1e-4)


def test_blending_image_of_the_reference_main_program():
    """BlendingRealRandomAccess.main (BlendingRealRandomAccess.java:251-270) renders a 500 x 500 image with border
    {100, 0} and blending {12, 150}.  The image is not stored in the reference, but it is fully determined: zero up to
    the border, the 1001-entry cosine table along the ramp, one on the plateau, product of the two axes, symmetric."""
    n = 500
    yy, xx = np.meshgrid(np.arange(n, dtype=np.float32), np.arange(n, dtype=np.float32), indexing="ij")
    t = np.stack([xx, yy, np.ones_like(xx)], axis=-1)
    # a third axis that never blends (negative border, range 1) turns the 3-d restatement into the 2-d program
    w = o.blending_weight(t, (n, n, 3), (100.0, 0.0, -10.0), (12.0, 150.0, 1.0))
    lut = o.blending_lookup()
    assert (w[:, :101] == 0).all() and (w[:, 399:] == 0).all() and (w[0] == 0).all() and (w[-1] == 0).all()
    assert (w[150:350, 112:388] == 1).all()                                  # plateau: 12 px behind the x border, 150 px in y
    np.testing.assert_array_equal(w, w[::-1, :])                             # symmetric in both axes
    np.testing.assert_array_equal(w, w[:, ::-1])
    for d in range(1, 12):                                                   # the ramp along x on a plateau row
        assert w[200, 100 + d] == np.float32(lut[int(np.floor(np.float32(d) / np.float32(12.0) * 1000.0 + 0.5))])
    for d in (1, 30, 75, 149):                                               # ... along y on a plateau column
        assert w[d, 250] == np.float32(lut[int(np.floor(np.float32(d) / np.float32(150.0) * 1000.0 + 0.5))])
    assert w[30, 105] == np.float32(np.float64(np.float32(lut[int(np.floor(np.float32(5) / np.float32(12.0) * 1000.0 + 0.5))])) *
                                    lut[int(np.floor(np.float32(30) / np.float32(150.0) * 1000.0 + 0.5))])   # product of the axes
