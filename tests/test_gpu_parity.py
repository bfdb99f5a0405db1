"""GPU parity tests: CUDA path through the C ABI vs the CPU oracle (oracle/oracle_np.py).

Tolerance: relative L2 <= 1e-4 after the configured iterations (BASELINE.json north_star);
single convolutions are held to 2e-6.  Integer work (tiling, copy/paste maps) is bit-exact.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_l2

pytestmark = pytest.mark.gpu

CONV_TOL = 2e-6
LOOP_TOL = 1e-4
# Statistics of an update step (sum and maximum of |psi' - psi|, MVDeconvolution.java:432-436).  A voxel's change is a
# difference of two float32 values of size psi, i.e. it carries an absolute error of ~1e-7 psi from the fp32 FFT, which
# is ~1e-4 of a typical change of 1e-3 psi: the sum averages that out, the maximum is a single voxel and keeps it.
# Observed on B200 over 10 - 30 iterations at K = 31^3 (profiles/r02_parity_growth.json): sum <= 2.4e-6, max <= 6.8e-6.
# Row partials of <= 32 voxels are summed in float before they enter the double sum: <= 32 * 2^-24 = 2e-6 relative.
STAT_SUM_TOL = 2e-5
STAT_MAX_TOL = 1e-4


@pytest.fixture(scope="module")
def m():
    import spim_registration_b200 as mod
    mod.load()
    devs = mod.CUDATools.queryCUDADetails()
    assert devs, "no CUDA device"
    return mod


@pytest.fixture(scope="module")
def o():
    import oracle_np
    return oracle_np


def _rand(shape, seed):
    return (np.random.default_rng(seed).random(shape, dtype=np.float32) + np.float32(0.1))


def _kernel(shape, seed):
    k = np.random.default_rng(seed).random(shape, dtype=np.float32)
    return (k / k.sum()).astype(np.float32)


@pytest.mark.parametrize("shape,ks", [
    ((20, 24, 28), (3, 5, 7)),       # smaller than one window
    ((33, 37, 41), (4, 6, 8)),       # ragged, even kernel
    ((5, 6, 7), (15, 15, 15)),       # volume smaller than the kernel halo: repeated reflection
    ((1, 40, 40), (1, 9, 9)),        # degenerate z
    ((64, 64, 64), (15, 15, 15)),
    ((70, 130, 150), (15, 11, 9)),   # several windows per axis, mixed window sizes
    ((96, 200, 260), (31, 31, 31)),
    ((200, 70, 150), (31, 9, 15)),   # z = one 256-point window (the warp-local z pass's second size)
    ((300, 36, 40), (31, 5, 7)),     # z needs the 512-point window
    ((530, 20, 24), (15, 3, 3)),     # z = 512-point window + a short one
    ((80, 150, 170), (63, 63, 63)),  # transformed PSFs grow with the view's affine (ExtractPSF.java:331-345)
    ((120, 60, 200), (95, 31, 95)),
    ((20, 40, 300), (5, 7, 151)),    # wider than a 128-point window allows: 256-point windows along x
    ((24, 280, 60), (3, 201, 9)),    # ... along y
    ((300, 30, 40), (255, 3, 5)),    # ... 512 points along z
])
@pytest.mark.parametrize("mode", ["zero", "const", "mirror"])
def test_convolution_matches_oracle(m, o, shape, ks, mode):
    f, k = _rand(shape, 1), _kernel(ks, 2)
    ref = o.convolve_extended(f, k, mode, 1.0)
    got = m.convolve(f, k, mode, 1.0)
    assert rel_l2(got, ref) <= CONV_TOL


def test_delta_lands_kernel_unflipped(m):
    """conv, not correlation (SURVEY 8c caveat): a delta at p puts kernel[j] at p + j - K//2."""
    k = np.arange(1, 3 * 5 * 7 + 1, dtype=np.float32).reshape(3, 5, 7)
    d = np.zeros((21, 23, 25), np.float32)
    d[10, 11, 12] = 1
    got = m.convolve(d, k, "zero")
    np.testing.assert_allclose(got[10 - 1:10 + 2, 11 - 2:11 + 3, 12 - 3:12 + 4], k, rtol=0, atol=2e-4)
    got[10 - 1:10 + 2, 11 - 2:11 + 3, 12 - 3:12 + 4] = 0
    assert np.abs(got).max() < 2e-4


def test_linearity(m):
    f, g, k = _rand((40, 44, 48), 3), _rand((40, 44, 48), 4), _kernel((7, 7, 7), 5)
    a = m.convolve(f + 2 * g, k, "zero")
    b = m.convolve(f, k, "zero") + 2 * m.convolve(g, k, "zero")
    assert rel_l2(a, b) <= CONV_TOL


def test_legacy_symbol_blockwise_equals_whole_volume(m, o):
    """Tiling invariance (SURVEY 8a invariant i): reference-style blocks through the drop-in
    symbol == whole-volume convolution, for both out-of-bounds rules."""
    f, k = _rand((50, 70, 90), 6), _kernel((9, 11, 13), 7)
    cuda = m.CUDAFourierConvolution()
    for mode, val in (("mirror", 0.0), ("const", 1.0)):
        whole = m.convolve(f, k, mode, val)
        blocks = m.BlockGeneratorFixedSizePrecise([64, 48, 40]).divideIntoBlocks(list(f.shape[::-1]), list(k.shape[::-1]))
        out = np.zeros_like(f)
        for b in blocks:
            data = o.copy_block(f, b.as_dict(), mode, val)  # Block.copyBlock from the extended source
            kk = k.copy()
            cuda.convolution3DfftCUDAInPlace(data, data.shape, kk, kk.shape, 0)  # dims (z,y,x) like the reference
            b.pasteBlock(out, data)
        assert rel_l2(out, whole) <= CONV_TOL
        assert rel_l2(out, o.convolve_extended(f, k, mode, val)) <= CONV_TOL


def test_legacy_device_queries(m, lib):
    n = lib.getNumDevicesCUDA()
    assert n >= 1
    assert lib.getCUDAcomputeCapabilityMajorVersion(0) == 10
    assert lib.getMemDeviceCUDA(0) > 100 * (1 << 30)
    assert 0 < lib.getFreeMemDeviceCUDA(0) <= lib.getMemDeviceCUDA(0)
    name = C.create_string_buffer(256)
    lib.getNameDeviceCUDA(0, name)
    assert b"B200" in name.value


def test_legacy_failure_leaves_block_untouched(m, lib):
    import spim_registration_b200._lib as L
    f = _rand((8, 8, 8), 8)
    before = f.copy()
    k = _kernel((3, 3, 3), 9)
    dim = (C.c_int * 3)(8, 8, 8)
    kd = (C.c_int * 3)(3, 3, 3)
    lib.convolution3DfftCUDAInPlace(f.ctypes.data_as(L.c_float_p), dim, k.ctypes.data_as(L.c_float_p), kd, 99)  # bad device
    code, msg = L.last_error()
    assert code != 0 and "device" in msg
    np.testing.assert_array_equal(f, before)
    kbig = np.ones((3, 3, 300), np.float32)  # x kernel wider than the largest xy window
    kd3 = (C.c_int * 3)(3, 3, 300)
    lib.convolution3DfftCUDAInPlace(f.ctypes.data_as(L.c_float_p), dim, kbig.ctypes.data_as(L.c_float_p), kd3, 0)
    code, msg = L.last_error()
    assert code != 0 and "window" in msg
    np.testing.assert_array_equal(f, before)


def _views(m, imgs, weights, psfs):
    inp = m.MVDeconInput()
    for img, w, p in zip(imgs, weights, psfs):
        inp.add(m.MVDeconFFT(img, w, p))
    return inp


def _dataset(o, shape, V, ksize, seed=0):
    from spim_registration_b200 import synthetic
    return synthetic.make_dataset(shape, V, ksize, seed=synthetic.SEED0 + seed, convolve=lambda f, k: o.convolve_extended(f, k, "mirror"))


@pytest.mark.parametrize("psftype", ["EFFICIENT_BAYESIAN", "OPTIMIZATION_I", "OPTIMIZATION_II", "INDEPENDENT"])
def test_kernel_preparation_matches_oracle(m, o, psftype):
    from spim_registration_b200 import synthetic
    psfs = synthetic.make_psfs(3, (9, 7, 11), extracted=True)
    inp = _views(m, [np.ones((4, 4, 4), np.float32)] * 3, [1.0] * 3, psfs)
    inp.init(psftype, 8)
    k1, k2 = o.init_kernels(psfs, psftype, 8)
    for v, view in enumerate(inp.getViews()):
        np.testing.assert_array_equal(view.getKernel1(), k1[v])  # host arithmetic only: bit-exact
        assert rel_l2(view.getKernel2(), k2[v]) <= 5e-6


@pytest.mark.parametrize("psftype,lam,use_weights", [
    ("EFFICIENT_BAYESIAN", 0.006, False),
    ("OPTIMIZATION_I", 0.006, True),
    ("OPTIMIZATION_II", 0.0, False),
    ("INDEPENDENT", 0.006, True),
])
def test_loop_matches_oracle(m, o, psftype, lam, use_weights):
    from spim_registration_b200 import synthetic
    shape, V, iters = (40, 52, 60), 3, 5
    ds = _dataset(o, shape, V, (9, 9, 9))
    weights = synthetic.blending_weights(shape, V) if use_weights else [np.float32(1.0 / V)] * V
    k1, k2 = o.init_kernels(ds["psfs"], psftype, 8)
    ref, stats, avg = o.run_deconvolution(ds["imgs"], weights, k1, k2, iters, lam)
    inp = _views(m, ds["imgs"], [w if use_weights else float(w) for w in weights], ds["psfs"])
    dec = m.MVDeconvolution(inp, psftype, iters, lam)
    assert abs(dec.avg - avg) <= 1e-12 * abs(avg)
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL
    rs = np.array([[s for s, _ in row] for row in stats])
    rm = np.array([[x for _, x in row] for row in stats])
    np.testing.assert_allclose(dec.sumChange, rs, rtol=STAT_SUM_TOL)
    np.testing.assert_allclose(dec.maxChange, rm, rtol=STAT_MAX_TOL)
    assert (dec.getPsi()[ref == 0] == 0).all()  # never-updated voxels masked (MVDeconvolution.java:180-187)


# ---- the named configurations' SETTINGS on oracle-feasible volumes (BASELINE.json configs[1..3]) -------------------
# The volumes span two overlap-save windows along every axis (112 = 128- + 64-point windows for K = 31 in x and y,
# 100 = 128 + 32 in z), the PSFs are 31^3 and the iteration counts are the configurations' own, so that the
# multi-window engine is compared with the oracle over 10 - 30 multiplicative iterations.  The relative L2 error is
# recorded after EVERY iteration (MVDeconvolution's debug stack, MVDeconvolution.java:139-175) -- fp32 FFT error
# does not accumulate: the update is a contraction towards the data, see profiles/r02_parity_growth.json.
CONFIG_SETTINGS = {
    "cfg2": dict(V=6, psftype="OPTIMIZATION_II", iters=20, extracted=True, kind="bead", dense_weights=False, direct=True),
    "cfg3": dict(V=7, psftype="EFFICIENT_BAYESIAN", iters=10, extracted=False, kind="blob", dense_weights=True, direct=True),
    "cfg3_fft_only": dict(V=7, psftype="EFFICIENT_BAYESIAN", iters=10, extracted=False, kind="blob", dense_weights=True, direct=False),
    "cfg4": dict(V=6, psftype="INDEPENDENT", iters=30, extracted=False, kind="blob", dense_weights=True, direct=True),
}


_ORACLE_RUNS = {}


@pytest.mark.parametrize("name", list(CONFIG_SETTINGS))
def test_loop_at_config_settings_matches_oracle_every_iteration(m, o, name):
    import json
    import os
    from spim_registration_b200 import synthetic
    cfg = CONFIG_SETTINGS[name]
    shape, V, iters = (100, 112, 112), cfg["V"], cfg["iters"]
    key = tuple(sorted((k, v) for k, v in cfg.items() if k != "direct"))
    if key not in _ORACLE_RUNS:   # the oracle's side is the same for the direct-path and the FFT-only variant
        truth = synthetic.make_truth(shape, cfg["kind"], synthetic.SEED0 + 2)
        psfs = synthetic.make_psfs(V, (31, 31, 31), extracted=cfg["extracted"], seed=synthetic.SEED0 + 2)
        imgs = [synthetic.degrade(o.convolve_extended(truth, (p.astype(np.float64) / p.astype(np.float64).sum()).astype(np.float32), "mirror"),
                                  v, V, synthetic.SEED0 + 2) for v, p in enumerate(psfs)]
        weights = synthetic.blending_weights(shape, V) if cfg["dense_weights"] else [np.float32(1.0 / V)] * V
        k1, k2 = o.init_kernels(psfs, cfg["psftype"], 8)
        ref_it = []
        ref, stats, avg = o.run_deconvolution(imgs, weights, k1, k2, iters, 0.006, on_iteration=lambda i, p: ref_it.append(p.copy()))
        _ORACLE_RUNS.clear()      # one entry at a time: 10 - 30 volumes each
        _ORACLE_RUNS[key] = (psfs, imgs, weights, ref_it, ref, stats)
    psfs, imgs, weights, ref_it, ref, stats = _ORACLE_RUNS[key]
    inp = _views(m, imgs, [w if cfg["dense_weights"] else float(w) for w in weights], psfs)
    dec = m.MVDeconvolution(inp, cfg["psftype"], iters, 0.006, debug=True, debugInterval=1, directPath=cfg["direct"])
    tiles = m.plan_tiles(shape[::-1], (31, 31, 31))
    for d in range(3):
        assert len({int(t[1][d]) for t in tiles}) >= 2, "volume must span two windows along every axis"
    if not cfg["direct"] or cfg["extracted"]:
        assert all(a == 0 and b == 0 for a, b in dec.directPath), dec.directPath   # everything on the FFT engine
    growth = [rel_l2(p, ref_it[i]) for i, p in dec.debugStack]
    assert len(growth) == iters
    rs = np.array([[s for s, _ in row] for row in stats])
    rm = np.array([[x for _, x in row] for row in stats])
    dsum = float(np.abs(dec.sumChange / rs - 1).max())
    dmax = float(np.abs(dec.maxChange / rm - 1).max())
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "parity_growth_%s.json" % name), "w") as fh:
            json.dump(dict(config=name, settings={k: v for k, v in cfg.items()}, volume_xyz=shape[::-1], kernel=31, rel_l2_per_iteration=growth,
                           sum_change_max_rel_dev=dsum, max_change_max_rel_dev=dmax, direct_path=dec.directPath), fh)
    assert max(growth) <= LOOP_TOL, growth
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL
    assert dsum <= STAT_SUM_TOL and dmax <= STAT_MAX_TOL, (dsum, dmax)


def test_golden_fixture(m):
    """CUDA path vs the frozen oracle output (tests/golden/make_golden.py)."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "loop_3view_32.npz"))
    V = int(g["num_views"])
    inp = m.MVDeconInput()
    for v in range(V):
        view = m.MVDeconFFT(g["img%d" % v], g["weight%d" % v], g["k1_%d" % v])
        view.kernel2 = g["k2_%d" % v]
        inp.add(view)
    dec = m.MVDeconvolution(inp, str(g["psftype"]), int(g["iterations"]), float(g["lam"]), prepared=True)
    assert rel_l2(dec.getPsi(), g["psi"]) <= LOOP_TOL


def test_fixed_point_independent(m, o):
    """SURVEY 8a invariant ii: with sum(kernel1) = 1 and img = conv1(truth), psi = truth is a
    fixed point of the INDEPENDENT update (lambda = 0, weight 1)."""
    from spim_registration_b200 import synthetic
    shape = (36, 40, 44)
    truth = synthetic.make_truth(shape, "blob", 5) + np.float32(0.05)
    k = _kernel((7, 7, 7), 11)
    img = o.convolve_extended(truth, k, "mirror")
    view = m.MVDeconFFT(img, 1.0, k)
    view.kernel2 = m.MVDeconFFT.computeInvertedKernel(k)
    inp = m.MVDeconInput()
    inp.add(view)
    dec = m.MVDeconvolution(inp, "INDEPENDENT", 3, 0.0, initialImage=truth, prepared=True)
    # conv2 of a quotient that is 1 inside is 1 only away from the border, where ext(q)=1 anyway
    assert rel_l2(dec.getPsi(), truth) <= 2e-5


def test_fixed_point_at_config3_size(m):
    """The fixed point of the INDEPENDENT update (SURVEY 8a invariant ii) at the 1024x1024x512 geometry of
    BASELINE.json configs[2..3]: two views with their own asymmetric 31^3 kernels, a dense weight image,
    a no-data slab.  img_v = conv1_v(truth) through the same engine, so the quotient is exactly 1 wherever
    there is data and psi must stay at truth in every voxel of every tile; the slab is zeroed by the final mask."""
    Z, Y, X = 512, 1024, 1024
    rng = np.random.default_rng(21)
    truth = rng.random((Z, Y, X), dtype=np.float32)
    truth += np.float32(0.05)
    weight = rng.random((Z, Y, X), dtype=np.float32)
    weight *= np.float32(0.7)
    weight += np.float32(0.3)
    inp = m.MVDeconInput()
    for v in range(2):
        k = _kernel((31, 31, 31), 30 + v)
        img = m.convolve(truth, k, "mirror")
        img[:, :, :100] = 0          # no data in either view
        view = m.MVDeconFFT(img, weight, k)
        view.kernel2 = m.MVDeconFFT.computeInvertedKernel(k)
        inp.add(view)
        del img
    dec = m.MVDeconvolution(inp, "INDEPENDENT", 1, 0.0, initialImage=truth, prepared=True)
    psi = dec.getPsi()
    assert (psi[:, :, :100] == 0).all()
    diff = psi[:, :, 100:] - truth[:, :, 100:]
    assert np.abs(diff).max() <= 1e-5
    assert rel_l2(psi[:, :, 100:], truth[:, :, 100:]) <= 2e-6


def test_no_data_and_constant_weight_edge_cases(m, o):
    shape = (24, 28, 36)
    rng = np.random.default_rng(3)
    imgs = [rng.random(shape, dtype=np.float32), np.zeros(shape, np.float32)]  # second view: no data at all
    imgs[0][:, :, :5] = 0
    psfs = [_kernel((5, 5, 5), 1), _kernel((5, 5, 5), 2)]
    k1, k2 = o.init_kernels(psfs, "INDEPENDENT", 8)
    ref, stats, avg = o.run_deconvolution(imgs, [np.float32(1.0), np.float32(0.3)], k1, k2, 3, 0.006)
    dec = m.MVDeconvolution(_views(m, imgs, [1.0, 0.3], psfs), "INDEPENDENT", 3, 0.006)
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL
    assert (dec.getPsi()[:, :, :5] == 0).all()


def test_config1_full_size(m, o):
    """BASELINE.json configs[0]: 3 views 128^3, Gaussian PSFs 15^3, efficient Bayesian, 10 iterations."""
    from spim_registration_b200 import synthetic
    shape, V, iters = (128, 128, 128), 3, 10
    ds = _dataset(o, shape, V, (15, 15, 15), seed=1)
    w = [np.float32(1.0 / V)] * V
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    ref, stats, avg = o.run_deconvolution(ds["imgs"], w, k1, k2, iters, 0.006)
    dec = m.MVDeconvolution(_views(m, ds["imgs"], [float(x) for x in w], ds["psfs"]), "EFFICIENT_BAYESIAN", iters, 0.006)
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL


def test_config3_full_size_convolution_properties(m):
    """BASELINE.json configs[2] geometry (1024x1024x512, kernel 31^3: 11 x 11 windows of 128 in xy, the 512- and
    64-point windows in z) checked through size-independent properties, since the oracle does not finish at this
    size: (i) a constant volume under mirror extension stays constant x sum(kernel) in every voxel of every tile;
    (ii) deltas on tile seams, volume faces and corners each land one unflipped copy of an asymmetric kernel
    (zero extension) and nothing else -- the volume sum equals the clipped kernel sums."""
    Z, Y, X = 512, 1024, 1024
    k = _kernel((31, 31, 31), 5)
    h = 15
    const = m.convolve(np.full((Z, Y, X), 0.75, np.float32), k, "mirror")
    want = np.float32(0.75) * np.float32(k.astype(np.float64).sum())
    assert np.abs(const - want).max() <= 4e-6 * want
    del const
    tiles = m.plan_tiles((X, Y, Z), (31, 31, 31))          # [n, (window, origin, first output voxel, output count), xyz]
    seam_x = sorted({int(t[2][0]) for t in tiles})[1]      # first output voxel of the second window along x
    seam_y = sorted({int(t[2][1]) for t in tiles})[2]
    seam_z = sorted({int(t[2][2]) for t in tiles})[-1]     # start of the short z window behind the 512-point one
    assert 0 < seam_x < X and 0 < seam_y < Y and 400 < seam_z < Z
    pts = [(0, 0, 0), (Z - 1, Y - 1, X - 1), (seam_z, seam_y, seam_x), (seam_z - 1, seam_y - 1, seam_x - 1), (200, 500, X - 1),
           (Z - 1, 40, 700), (300, seam_y, 3), (seam_z, 900, seam_x + 1)]
    f = np.zeros((Z, Y, X), np.float32)
    for i, p in enumerate(pts):
        f[p] = 1.0 + i
    out = m.convolve(f, k, "zero")
    expected_sum = 0.0
    for i, (z, y, x) in enumerate(pts):
        lo = [max(0, c - h) for c in (z, y, x)]
        hi = [min(n, c + h + 1) for c, n in zip((z, y, x), (Z, Y, X))]
        want_patch = np.zeros([b - a for a, b in zip(lo, hi)], np.float64)
        for j, q in enumerate(pts):   # neighbouring deltas overlap: every delta's footprint inside this patch
            a = [max(l, c - h) for l, c in zip(lo, q)]
            b = [min(u, c + h + 1) for u, c in zip(hi, q)]
            if all(bb > aa for aa, bb in zip(a, b)):
                want_patch[a[0] - lo[0]:b[0] - lo[0], a[1] - lo[1]:b[1] - lo[1], a[2] - lo[2]:b[2] - lo[2]] += (1.0 + j) * k[
                    a[0] - q[0] + h:b[0] - q[0] + h, a[1] - q[1] + h:b[1] - q[1] + h, a[2] - q[2] + h:b[2] - q[2] + h]
        patch = out[lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]]
        assert np.abs(patch - want_patch).max() <= 2e-6 * want_patch.max()
        own = k[lo[0] - z + h:hi[0] - z + h, lo[1] - y + h:hi[1] - y + h, lo[2] - x + h:hi[2] - x + h]
        expected_sum += (1.0 + i) * float(own.astype(np.float64).sum())
    assert abs(float(out.sum(dtype=np.float64)) - expected_sum) <= 1e-4 * expected_sum
    assert float(np.abs(out).sum(dtype=np.float64)) <= expected_sum * (1 + 1e-3)   # nothing of either sign elsewhere


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("grid,shape", [((2, 1, 1), (40, 56, 72)), ((1, 2, 2), (40, 56, 72)), ((2, 2, 2), (40, 56, 72)),
                                        ((2, 2, 2), (41, 54, 70)), ((4, 1, 2), (40, 30, 150))])
def test_bricks_with_ghost_layers_equal_whole_volume(m, o, grid, shape, fused):
    """The brick / ghost-layer code paths of the library (what every rank of the multi-GPU run
    executes) on one device: all bricks stepped in lock-step must reproduce the single-context
    result and the oracle -- with in-process ghost copies between the steps, and with the fused
    exchange (mvdAttachPeer: the step kernels store into the neighbours' ghost layers)."""
    from spim_registration_b200 import synthetic
    from spim_registration_b200.parallel import LocalBrickGroup
    V, iters = 2, 2
    ds = _dataset(o, shape, V, (9, 7, 11), seed=3)
    weights = synthetic.blending_weights(shape, V)
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    ref, _, avg = o.run_deconvolution(ds["imgs"], weights, k1, k2, iters, 0.006)
    grp = LocalBrickGroup(shape[::-1], grid, ds["imgs"], weights, k1, k2, 0.006)
    try:
        assert abs(grp.init_psi_average() - avg) <= 1e-9 * abs(avg)
        grp._exchange(0)
        if fused:
            grp.attach_peers()
        grp.run_iterations(iters)
        got = grp.get_psi()
    finally:
        grp.close()
    assert rel_l2(got, ref) <= LOOP_TOL


@pytest.mark.parametrize("direct", [0, 1])
def test_fused_ghost_exchange_is_bit_identical_to_copies(m, o, direct):
    """Pushing ghost voxels from the epilogue changes where values are stored, not the values -- for the FFT
    engine (direct = 0) and for the direct separable kernels the Gaussian PSFs of this data set qualify for."""
    from spim_registration_b200 import synthetic
    from spim_registration_b200.parallel import LocalBrickGroup
    shape, V, grid = (37, 50, 150), 2, (2, 2, 2)
    ds = _dataset(o, shape, V, (9, 7, 11), seed=5)
    weights = synthetic.blending_weights(shape, V)
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    out = []
    for fused in (False, True):
        grp = LocalBrickGroup(shape[::-1], grid, ds["imgs"], weights, k1, k2, 0.006)
        try:
            for b in grp.bricks:
                b.L.check(b.lib.mvdSetDirectPath(b.ctx, direct), "mvdSetDirectPath")
                a_, b_ = C.c_int(), C.c_int()
                b.L.check(b.lib.mvdViewUsesDirectPath(b.ctx, 0, C.byref(a_), C.byref(b_)), "mvdViewUsesDirectPath")
                assert bool(a_.value) == bool(direct) and bool(b_.value) == bool(direct)
            grp.init_psi_average()
            if fused:
                grp.attach_peers()
            grp.run_iterations(3)
            out.append(grp.get_psi())
        finally:
            grp.close()
    np.testing.assert_array_equal(out[0], out[1])


# ---- direct separable path and the separable-convolution hook -------------------------------------
@pytest.mark.parametrize("size,shape", [(7, (20, 33, 150)), (15, (40, 70, 260)), (31, (35, 66, 140)), (63, (70, 80, 90)), (127, (140, 150, 160))])
@pytest.mark.parametrize("oob", [0, 1, 2])
def test_separable_hook_matches_oracle(m, o, size, shape, oob):
    """convolve_N (CUDASeparableConvolution.java:13-17; convolve_127 is called at CUDASeparableConvolutionFunctions.java:191):
    N <= 31 direct kernels, 63 and 127 through the FFT engine axis by axis."""
    rng = np.random.default_rng(size + oob)
    f = rng.random(shape, dtype=np.float32) + np.float32(0.1)
    ks = [np.zeros(size, np.float32) for _ in range(3)]
    for i, n in enumerate((size, max(3, size - 4), max(3, size // 2 | 1))):  # kernels shorter than N are centred (getFloatKernelPadded)
        k = rng.random(n).astype(np.float32)
        k /= k.sum()
        ks[i][(size - n) // 2:(size - n) // 2 + n] = k
    ref = o.separable_convolve(f, ks[0], ks[1], ks[2], ("zero", "const", "clamp")[oob], 0.6)
    got = f.copy()
    cuda = m.CUDASeparableConvolution()
    assert cuda.convolve(size, got, ks[0], ks[1], ks[2], shape[2], shape[1], shape[0], True, True, True, oob, 0.6, 0)
    assert rel_l2(got, ref) <= CONV_TOL


def test_separable_hook_partial_axes_gauss_and_refusals(m, o):
    F = m.CUDASeparableConvolutionFunctions
    rng = np.random.default_rng(9)
    f = rng.random((30, 40, 50), dtype=np.float32)
    k = o.cuda_kernels([2.0])[0]
    cuda = m.CUDASeparableConvolution()
    got = f.copy()
    assert cuda.convolve(len(k), got, None, k, None, 50, 40, 30, False, True, False, 2, 0.0, 0)  # y only
    assert rel_l2(got, o.separable_convolve(f, None, k, None, "clamp")) <= CONV_TOL
    # gauss(): kernels from sigma, extend-border rule (CUDASeparableConvolutionFunctions.java:33-58, 127-197)
    got = f.copy()
    assert F.gauss(got, [1.8, 1.8, 2.5])
    ks = o.cuda_kernels([1.8, 1.8, 2.5])
    assert rel_l2(got, o.separable_convolve(f, ks[0], ks[1], ks[2], "clamp")) <= CONV_TOL
    img2 = rng.random((64, 72), dtype=np.float32)  # 2-d image: z kernel null, d = 1
    got2 = img2.copy()
    assert F.gauss(got2, [3.0, 1.2])
    k2 = o.cuda_kernels([3.0, 1.2])
    assert rel_l2(got2, o.separable_convolve(img2[None], k2[0], k2[1], None, "clamp")[0]) <= CONV_TOL
    # refusals leave the image untouched and return false
    before = f.copy()
    assert not cuda.convolve(7, f, k[:7].copy(), None, None, 50, 40, 30, True, False, False, 5, 0.0, 0)  # bad out-of-bounds code
    assert not cuda.convolve(7, f, k[:7].copy(), None, None, 50, 40, 30, True, False, False, 0, 0.0, 99)  # bad device
    np.testing.assert_array_equal(f, before)


@pytest.mark.parametrize("psftype,ksize", [("EFFICIENT_BAYESIAN", (9, 9, 9)), ("OPTIMIZATION_I", (15, 11, 7)), ("INDEPENDENT", (5, 7, 9))])
def test_direct_path_for_gaussian_psfs_matches_oracle_and_fft_path(m, o, psftype, ksize):
    """Gaussian PSFs are rank-1: conv1 (and the compound kernels of conv2 when they stay within 31 taps)
    run as direct separable passes -- same result as the oracle and as the FFT engine."""
    from spim_registration_b200 import synthetic
    shape, V, iters = (40, 52, 140), 3, 4

    def g(n, sigma):
        x = np.arange(n, dtype=np.float64) - n // 2
        return np.exp(-x * x / (2 * sigma * sigma))
    psfs = [(g(ksize[0], 1.0 + 0.4 * v)[:, None, None] * g(ksize[1], 1.6 - 0.3 * v)[None, :, None] *
             g(ksize[2], 1.2 + 0.2 * v)[None, None, :]).astype(np.float32) for v in range(V)]  # axis-aligned: rank 1
    truth = synthetic.make_truth(shape, "blob", synthetic.SEED0 + 7)
    imgs = [synthetic.degrade(o.convolve_extended(truth, (p / p.sum()).astype(np.float32), "mirror"), v, V, synthetic.SEED0 + 7)
            for v, p in enumerate(psfs)]
    ds = dict(psfs=psfs, imgs=imgs)
    weights = synthetic.blending_weights(shape, V)
    k1, k2 = o.init_kernels(ds["psfs"], psftype, 8)
    ref, stats, _ = o.run_deconvolution(ds["imgs"], weights, k1, k2, iters, 0.006)
    res = {}
    for direct in (True, False):
        dec = m.MVDeconvolution(_views(m, ds["imgs"], weights, ds["psfs"]), psftype, iters, 0.006, directPath=direct)
        res[direct] = dec
        assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL
    assert all(a > 0 for a, _ in res[True].directPath), res[True].directPath   # every conv1 is direct
    assert all(a == 0 and b == 0 for a, b in res[False].directPath)
    assert rel_l2(res[True].getPsi(), res[False].getPsi()) <= LOOP_TOL
    rs = np.array([[s for s, _ in row] for row in stats])
    np.testing.assert_allclose(res[True].sumChange, rs, rtol=STAT_SUM_TOL)


def test_extracted_psfs_stay_on_the_fft_engine(m, o):
    """Bead-extracted PSFs (noise, rotation) are not rank 1: every convolution goes through the FFT engine."""
    from spim_registration_b200 import synthetic
    psfs = synthetic.make_psfs(2, (9, 9, 9), extracted=True)
    imgs = [_rand((24, 28, 32), 20 + v) for v in range(2)]
    dec = m.MVDeconvolution(_views(m, imgs, [0.5, 0.5], psfs), "EFFICIENT_BAYESIAN", 1, 0.006)
    assert all(a == 0 and b == 0 for a, b in dec.directPath), dec.directPath


# ---- inputs of the loop produced on the device ---------------------------------
def _registration(angle_deg, scale_z, shift):
    a = np.deg2rad(angle_deg)
    m = np.eye(4)
    m[:3, :3] = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]]) @ np.diag([1.0, 1.0, scale_z])
    m[:3, 3] = shift
    return m


@pytest.mark.parametrize("angle,border,blend", [(0.0, (0, 0, 0), (8, 8, 4)), (35.0, (3, 2, 1), (10, 12, 6)), (90.0, (-2, 0, 0), (5, 5, 5))])
def test_transform_view_and_weights_match_oracle(m, o, angle, border, blend):
    """mvdTransformView == TransformInputAndWeights restated by the oracle: resampled view bit-close, weights exact."""
    rng = np.random.default_rng(11)
    img = rng.random((30, 44, 52), dtype=np.float32) * np.float32(0.9)
    reg = _registration(angle, 2.5, (20.0, 5.0, 30.0))
    offset, dims = (-10, -4, -8), (120, 60, 110)
    got_t, got_w = m.transformInputAndWeights(img, reg, offset, dims, border, blend)
    inv = np.linalg.inv(reg)[:3, :4]
    ref_t, ref_w = o.transform_input_and_weights(img, inv, offset, dims, border, blend)
    assert (ref_t > 0).mean() > 0.05                       # the view really lies inside the box
    np.testing.assert_array_equal(got_t == 0, ref_t == 0)  # same "no data" set
    np.testing.assert_allclose(got_t, ref_t, rtol=0, atol=2e-7)
    np.testing.assert_allclose(got_w, ref_w, rtol=0, atol=1e-7)
    only_t, none_w = m.transformInputAndWeights(img, reg, offset, dims)  # TransformInput: no weights
    assert none_w is None
    np.testing.assert_array_equal(only_t, got_t)


def test_weight_normalizer_matches_oracle(m, o):
    rng = np.random.default_rng(12)
    ws = [np.where(rng.random((20, 30, 40)) < 0.3, 0, rng.random((20, 30, 40))).astype(np.float32) for _ in range(4)]
    ref, mn, avg = o.normalize_weights(ws)
    for speed in (1.0, 2.0):
        wn = m.WeightNormalizer([w.copy() for w in ws])
        assert wn.process(speed)
        assert wn.getMinOverlappingViews() == mn and abs(wn.getAvgOverlappingViews() - avg) < 1e-9
        for a, b in zip(wn.weights, ref):
            np.testing.assert_array_equal(a, o.adjust_for_osem(b, speed))  # NaN == NaN position-wise, values bit-exact


@pytest.mark.parametrize("osem", [1.0, 2.0])
def test_virtual_weights_match_oracle(m, o, osem):
    """WeightType.VIRTUAL_WEIGHTS materialised on the device == NormalizingRandomAccess over the blending functions
    restated by the oracle: sum image, overlap statistics and every view's weights bit-exact."""
    offset, dims = (-10, -4, -8), (120, 60, 110)
    views = [((52, 44, 30), _registration(0.0, 2.5, (20.0, 5.0, 30.0)), (0, 0, 0), (8, 8, 4)),
             ((52, 44, 30), _registration(35.0, 2.5, (25.0, 3.0, 20.0)), (3, 2, 1), (10, 12, 6)),
             ((40, 50, 36), _registration(90.0, 2.0, (60.0, 0.0, 10.0)), (-2, 0, 0), (5, 5, 5))]
    vw = m.VirtualWeights(views, offset, dims)
    assert vw.process()
    blends = [o.virtual_blending(np.linalg.inv(reg)[:3, :4], offset, dims, d, b, r) for d, reg, b, r in views]
    ref_sum, mn, avg = o.compute_sum_image(blends)
    assert (ref_sum > 1).mean() > 0.01 and (ref_sum == 1).mean() > 0.01   # both branches of the sum image
    np.testing.assert_array_equal(vw.sumWeights, ref_sum)
    assert vw.minOverlappingViews == mn and abs(vw.avgOverlappingViews - avg) < 1e-9
    vw.setOSEMspeedup(osem)
    for v in range(len(views)):
        np.testing.assert_array_equal(vw.weight(v), o.normalizing_weight(blends[v], ref_sum, osem))


def test_virtual_weights_equal_the_blending_of_transform_view(m, o):
    """The two routes to a blending weight (TransformInputAndWeights and the virtual accessor) see the same positions."""
    reg = _registration(35.0, 2.5, (20.0, 5.0, 30.0))
    offset, dims = (-10, -4, -8), (96, 60, 70)
    img = np.ones((30, 44, 52), np.float32)
    _, w = m.transformInputAndWeights(img, reg, offset, dims, (3, 2, 1), (10, 12, 6))
    vw = m.VirtualWeights([((52, 44, 30), reg, (3, 2, 1), (10, 12, 6))], offset, dims)
    vw.process()
    np.testing.assert_array_equal(vw.weight(0), o.normalizing_weight(w, vw.sumWeights))


def test_device_prepared_inputs_drive_the_loop(m, o):
    """Views resampled and weighted on the device, then deconvolved: same result as the oracle fed with the oracle's inputs."""
    from spim_registration_b200 import synthetic
    rng = np.random.default_rng(13)
    dims, offset, V = (64, 48, 40), (0, 0, 0), 3
    truth = synthetic.make_truth(dims[::-1], "blob", 5)
    imgs_d, ws_d, imgs_o, ws_o = [], [], [], []
    for v in range(V):
        reg = _registration(0.0, 1.0, (3.0 * v, -2.0 * v, 1.5 * v))
        raw = (truth + np.float32(0.02) * rng.random(truth.shape, dtype=np.float32)).astype(np.float32)
        t, w = m.transformInputAndWeights(raw, reg, offset, dims, (0, 0, 0), (6, 6, 6))
        to, wo = o.transform_input_and_weights(raw, np.linalg.inv(reg)[:3, :4], offset, dims, (0, 0, 0), (6, 6, 6))
        imgs_d.append(t); ws_d.append(w); imgs_o.append(to); ws_o.append(wo)
    wn = m.WeightNormalizer(ws_d)
    wn.process(1.0)
    ws_o, _, _ = o.normalize_weights(ws_o)
    ws_dn = [np.nan_to_num(w, nan=0.0) for w in wn.weights]   # voxels outside every view: no data, weight irrelevant
    ws_on = [np.nan_to_num(w, nan=0.0) for w in ws_o]
    psfs = synthetic.make_psfs(V, (7, 7, 7), extracted=True)
    k1, k2 = o.init_kernels(psfs, "EFFICIENT_BAYESIAN", 8)
    ref, _, _ = o.run_deconvolution(imgs_o, ws_on, k1, k2, 3, 0.006)
    dec = m.MVDeconvolution(_views(m, imgs_d, ws_dn, psfs), "EFFICIENT_BAYESIAN", 3, 0.006)
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL


# ---- edge cases of the loop ------------------------------------------------------------------------
@pytest.mark.parametrize("shape,ksize", [
    ((1, 40, 44), (1, 7, 7)),        # a single plane: 2-d deconvolution
    ((9, 10, 11), (15, 15, 15)),     # kernel larger than the volume: repeated reflection in conv1, mostly "outside = 1" in conv2
    ((33, 37, 41), (5, 6, 8)),       # odd sizes, even kernel extents (centre K // 2, Mirror's even-size quirk in kernel2)
    ((20, 24, 131), (3, 3, 31)),     # widest kernel the 128-wide window supports, in x only
])
def test_loop_edge_shapes_match_oracle(m, o, shape, ksize):
    rng = np.random.default_rng(21)
    V, iters = 2, 3
    psfs = [(rng.random(ksize, dtype=np.float32) + np.float32(0.05)) for _ in range(V)]
    imgs = [rng.random(shape, dtype=np.float32) + np.float32(0.05) for _ in range(V)]
    imgs[1][..., : shape[2] // 3] = 0                                  # a slab without data in one view
    weights = [np.float32(0.5)] * V
    k1, k2 = o.init_kernels(psfs, "OPTIMIZATION_I", 8)
    ref, _, _ = o.run_deconvolution(imgs, weights, k1, k2, iters, 0.006)
    dec = m.MVDeconvolution(_views(m, imgs, [0.5, 0.5], psfs), "OPTIMIZATION_I", iters, 0.006)
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL


def test_view_without_any_data_and_never_updated_voxels(m, o):
    """One view is empty (all zeros = no data): its quotient is 1 everywhere and it must not disturb the result;
    voxels no view covers end as 0 (MVDeconvolution.java:180-187)."""
    rng = np.random.default_rng(22)
    shape = (24, 28, 36)
    psfs = [(rng.random((5, 5, 5), dtype=np.float32) + np.float32(0.05)) for _ in range(3)]
    imgs = [rng.random(shape, dtype=np.float32) + np.float32(0.05) for _ in range(3)]
    imgs[2][:] = 0
    imgs[0][:, :, 30:] = 0
    imgs[1][:, :, 30:] = 0
    k1, k2 = o.init_kernels(psfs, "EFFICIENT_BAYESIAN", 8)
    ref, _, _ = o.run_deconvolution(imgs, [np.float32(1 / 3)] * 3, k1, k2, 3, 0.006)
    dec = m.MVDeconvolution(_views(m, imgs, [1 / 3] * 3, psfs), "EFFICIENT_BAYESIAN", 3, 0.006)
    assert rel_l2(dec.getPsi(), ref) <= LOOP_TOL
    assert (dec.getPsi()[:, :, 30:] == 0).all() and (dec.getPsi()[:, :, :30] > 0).all()


# ---- PSF extraction on the device ------------------------------------------------
def test_extract_and_transform_psf_match_oracle(m, o):
    rng = np.random.default_rng(41)
    img = rng.random((40, 60, 70), dtype=np.float32)
    locs = np.column_stack([rng.uniform(0, 70, 25), rng.uniform(0, 60, 25), rng.uniform(0, 40, 25)])
    locs[0] = (0.25, 59.5, 39.75)                                    # windows that wrap around (periodic extension)
    e = m.ExtractPSF()
    size = (19, 17, 11)
    got = e.extractPSFLocal(img, locs, size)
    ref = o.extract_psf_local(img, locs, size)
    np.testing.assert_allclose(got, ref, rtol=0, atol=4e-6)         # 25 float additions of values < 1
    gotn = e.extractPSFLocal(img, locs, size, normalize=True)
    np.testing.assert_allclose(gotn, o.normalize_psf(ref), rtol=0, atol=2e-6)
    for angle, sz in ((0.0, 1.0), (30.0, 2.5), (90.0, 3.0)):
        model = _registration(angle, sz, (12.0, -3.0, 7.5))
        t = e.transformPSF(gotn, model)
        rt = o.transform_psf(gotn, model[:3, :4])
        assert t.shape == rt.shape and all(s % 2 == 1 for s in t.shape)
        np.testing.assert_allclose(t, rt, rtol=0, atol=2e-7)
    psf = e.extractNextImg(img, _registration(45.0, 2.0, (0, 0, 0)), locs, size)
    assert psf.shape == e.pointSpreadFunctions[0].shape and len(e.originalPSFs) == 1


# ---- several devices behind one handle of the C ABI (mvdCreateMulti), one process ------------------------------------
def _multi_dataset(o):
    from spim_registration_b200 import synthetic
    shape, V = (44, 56, 72), 3
    ds = _dataset(o, shape, V, (9, 7, 11), seed=9)
    weights = synthetic.blending_weights(shape, V)
    return shape, V, ds, weights


@pytest.mark.parametrize("ndev", [2, 4, 8])
def test_multi_device_handle_on_one_gpu_equals_single_context(m, o, ndev):
    """mvdCreateMulti with the same device named several times: bricks, peer attachment, the per-device worker threads
    and the event-based neighbour barrier are exactly what several GPUs run (MVDeconFFT.java:426-445 drives its devices
    from one process too); the result must equal the oracle and the single-context run, statistics included."""
    shape, V, ds, weights = _multi_dataset(o)
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    ref, stats, avg = o.run_deconvolution(ds["imgs"], weights, k1, k2, 3, 0.006)
    single = m.MVDeconvolution(_views(m, ds["imgs"], weights, ds["psfs"]), "EFFICIENT_BAYESIAN", 3, 0.006)
    inp = m.MVDeconInput()
    for img, w, p in zip(ds["imgs"], weights, ds["psfs"]):
        inp.add(m.MVDeconFFT(img, w, p, deviceList=[0] * ndev))
    multi = m.MVDeconvolution(inp, "EFFICIENT_BAYESIAN", 3, 0.006)
    assert len(multi.bricks) == ndev
    assert abs(multi.avg - avg) <= 1e-9 * abs(avg)
    assert rel_l2(multi.getPsi(), ref) <= LOOP_TOL
    assert rel_l2(multi.getPsi(), single.getPsi()) <= 2e-6
    rs = np.array([[s for s, _ in row] for row in stats])
    np.testing.assert_allclose(multi.sumChange, rs, rtol=STAT_SUM_TOL)
    np.testing.assert_allclose(multi.maxChange, np.array([[x for _, x in row] for row in stats]), rtol=STAT_MAX_TOL)


def test_multi_device_handle_initial_image_and_debug_stack(m, o):
    """mvdMultiSetPsi fills the bricks' ghost layers from the caller's volume; stepping one iteration at a time
    (debug stack) keeps the neighbour events consistent across calls."""
    shape, V, ds, weights = _multi_dataset(o)
    k1, k2 = o.init_kernels(ds["psfs"], "INDEPENDENT", 8)
    psi0 = (np.random.default_rng(4).random(shape, dtype=np.float32) + np.float32(0.2))
    ref_it = []
    ref, _, _ = o.run_deconvolution(ds["imgs"], weights, k1, k2, 3, 0.006, psi0=psi0, on_iteration=lambda i, p: ref_it.append(p.copy()))
    inp = m.MVDeconInput()
    for img, w, p in zip(ds["imgs"], weights, ds["psfs"]):
        inp.add(m.MVDeconFFT(img, w, p, deviceList=[0, 0, 0, 0]))
    multi = m.MVDeconvolution(inp, "INDEPENDENT", 3, 0.006, initialImage=psi0, debug=True)
    for it, snap in multi.debugStack:
        assert rel_l2(snap, ref_it[it]) <= LOOP_TOL
    assert rel_l2(multi.getPsi(), ref) <= LOOP_TOL


def test_multi_device_handle_on_real_devices(m, o, lib):
    """The same through 2 (4, 8) distinct GPUs when the box has them; skipped on a single-GPU box."""
    n = lib.getNumDevicesCUDA()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    ndev = 8 if n >= 8 else (4 if n >= 4 else 2)
    shape, V, ds, weights = _multi_dataset(o)
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    ref, stats, avg = o.run_deconvolution(ds["imgs"], weights, k1, k2, 3, 0.006)
    inp = m.MVDeconInput()
    for img, w, p in zip(ds["imgs"], weights, ds["psfs"]):
        inp.add(m.MVDeconFFT(img, w, p, deviceList=list(range(ndev))))
    multi = m.MVDeconvolution(inp, "EFFICIENT_BAYESIAN", 3, 0.006)
    assert sorted(b[0] for b in multi.bricks) == list(range(ndev))
    assert rel_l2(multi.getPsi(), ref) <= LOOP_TOL


# ---- the TMA kernel generations (off by default: measured slower) stay correct ---------------------
@pytest.mark.parametrize("switch,value", [("z_tma", 1), ("inv_ring", 1), ("inv_ring", 0), ("z_warp", 0), ("z_warp", 1), ("z_warp", 2),
                                          ("z_warp", 3), ("z_warp", 4)])
def test_tma_kernels_match_oracle_and_default_kernels(m, o, lib, switch, value):
    """Every kernel generation against the default one and the oracle: the z pass with bulk-tensor loads / stores, the inverse
    xy ring kernel fed by bulk-tensor loads, and (z_warp = 0) the register-loading z pass against the default warp-local z
    pass.  The arithmetic is the same, only the data movement differs."""
    from spim_registration_b200 import synthetic
    zdim = 200 if (switch, value) in (("z_warp", 2), ("z_warp", 4)) else 300   # z_warp = 2 / 4: the 256-point z window
    f, k = _rand((zdim, 150, 260), 31), _kernel((31, 31, 31), 32)      # 512-point z window, two 128-windows in x and y
    ref = o.convolve_extended(f, k, "mirror")
    base = m.convolve(f, k, "mirror")
    shape, V = (140, 112, 120), 2
    ds = _dataset(o, shape, V, (15, 15, 15), seed=11)
    weights = synthetic.blending_weights(shape, V)
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    loop_ref, _, _ = o.run_deconvolution(ds["imgs"], weights, k1, k2, 2, 0.006)
    prev = lib.mvdSetTuning(switch.encode(), value)
    try:
        assert prev in (0, 1, 2, 3, 4)
        got = m.convolve(f, k, "mirror")
        dec = m.MVDeconvolution(_views(m, ds["imgs"], weights, ds["psfs"]), "EFFICIENT_BAYESIAN", 2, 0.006, directPath=False)
    finally:
        lib.mvdSetTuning(switch.encode(), prev)
    assert rel_l2(got, ref) <= CONV_TOL
    assert rel_l2(got, base) <= 1e-7
    assert rel_l2(dec.getPsi(), loop_ref) <= LOOP_TOL


# ---- known-answer programs of the reference through the DEVICE code ------------------------------------------------------
def test_tikhonov_table_through_the_device_epilogue(m, o):
    """MVDeconvolution.main (MVDeconvolution.java:707-714) prints tikhonov(d * 10000, 0.0006) and tikhonov(d, 0.0006)
    for d = 0, 0.1, ... < 10.  The same 200 values go through the update epilogue of the device: psi = 1, delta kernels
    (conv1(psi) = 1, conv2(q) = q exactly on the direct path), so value = img and psi' = 1 + (tikhonov(img) - 1)."""
    d = np.arange(0, 100, dtype=np.float64) * 0.1
    vals = np.concatenate([d * 10000.0, d]).astype(np.float32).reshape(1, 10, 20)
    lam = 0.0006
    one = np.ones((1, 1, 1), np.float32)
    view = m.MVDeconFFT(vals, 1.0, one)
    view.kernel2 = one.copy()
    inp = m.MVDeconInput()
    inp.add(view)
    psi0 = np.ones(vals.shape, np.float32)
    dec = m.MVDeconvolution(inp, "INDEPENDENT", 1, lam, initialImage=psi0, prepared=True)
    assert dec.directPath == [(7, 7)]
    integral = np.where(vals > 0, vals, np.float32(1))            # no data -> quotient 1
    want, _, _ = o.compute_final_values(psi0, integral, np.float32(1.0), lam)
    want = np.where(vals > 0, want, np.float32(0))                # never-updated voxels are masked
    got = dec.getPsi()
    np.testing.assert_allclose(got, want, rtol=1.5e-7, atol=0)    # <= 1 ulp: the epilogue evaluates the term in two-float arithmetic
    assert (got == want).mean() >= 0.95


def test_blending_image_of_the_reference_main_program_on_the_device(m, o):
    """BlendingRealRandomAccess.main (:251-270): 500 x 500, border {100, 0}, blending {12, 150} -- rendered by the device's
    blending code (mvdTransformView) and compared bit for bit with the restated image."""
    n = 500
    img = np.ones((3, n, n), np.float32)
    _, w = m.transformInputAndWeights(img, np.eye(4), (0, 0, 1), (n, n, 1), (100.0, 0.0, -10.0), (12.0, 150.0, 1.0))
    yy, xx = np.meshgrid(np.arange(n, dtype=np.float32), np.arange(n, dtype=np.float32), indexing="ij")
    ref = o.blending_weight(np.stack([xx, yy, np.ones_like(xx)], axis=-1), (n, n, 3), (100.0, 0.0, -10.0), (12.0, 150.0, 1.0))
    np.testing.assert_array_equal(w[0], ref)
    assert (w[0][150:350, 112:388] == 1).all() and (w[0][:, :101] == 0).all()


def test_direct_path_near_its_threshold_and_small_lambda(m, o):
    """the rank-1 substitution (sum |K - a x b x c| <= 1e-6 sum |K|) is an approximation the reference does
    not make; hold it at the edge of its acceptance threshold, with a small lambda (where a cancelling Tikhonov evaluation
    would lose bits), over 20 iterations: same result as the oracle on the kernel AS GIVEN and as the FFT engine."""
    from spim_registration_b200 import synthetic
    shape, V, iters, lam = (40, 48, 72), 2, 20, 0.0006
    rng = np.random.default_rng(5)

    def g(n, sigma):
        x = np.arange(n, dtype=np.float64) - n // 2
        return np.exp(-x * x / (2 * sigma * sigma))
    psfs = []
    for v in range(V):
        sep = g(9, 1.3 + 0.3 * v)[:, None, None] * g(9, 1.1)[None, :, None] * g(9, 1.6 - 0.2 * v)[None, None, :]
        noise = rng.standard_normal(sep.shape)
        # a perturbation of 5e-7 of the kernel's L1 norm: still accepted as rank 1 after float rounding and normalisation
        psfs.append((sep + noise * (5e-7 * np.abs(sep).sum() / np.abs(noise).sum())).astype(np.float32))
    truth = synthetic.make_truth(shape, "blob", synthetic.SEED0 + 9)
    imgs = [synthetic.degrade(o.convolve_extended(truth, (p / p.sum()).astype(np.float32), "mirror"), v, V, synthetic.SEED0 + 9)
            for v, p in enumerate(psfs)]
    weights = [np.float32(0.5)] * V
    k1, k2 = o.init_kernels(psfs, "INDEPENDENT", 8)
    ref, stats, _ = o.run_deconvolution(imgs, weights, k1, k2, iters, lam)
    res = {}
    for direct in (True, False):
        res[direct] = m.MVDeconvolution(_views(m, imgs, [0.5] * V, psfs), "INDEPENDENT", iters, lam, directPath=direct)
        assert rel_l2(res[direct].getPsi(), ref) <= LOOP_TOL
    assert all(a > 0 and b > 0 for a, b in res[True].directPath), res[True].directPath
    assert all(a == 0 and b == 0 for a, b in res[False].directPath)
    assert rel_l2(res[True].getPsi(), res[False].getPsi()) <= 2e-5
    rs = np.array([[s for s, _ in row] for row in stats])
    # the substituted kernel differs by ~1e-6 of its norm: changes of ~1e-3 psi move by up to ~1e-3 of themselves
    np.testing.assert_allclose(res[True].sumChange, rs, rtol=2e-3)


# ---- kernel preparation behind the C ABI and the declared-but-unused convolutionCPU ----------------------
@pytest.mark.parametrize("psftype", ["EFFICIENT_BAYESIAN", "OPTIMIZATION_I", "OPTIMIZATION_II", "INDEPENDENT"])
def test_prepare_kernels_entry_matches_oracle_and_host_mirror(m, o, psftype):
    """mvdPrepareKernels == MVDeconInput.init: kernel1 bit-exact against the oracle (host arithmetic only, sumImg quirk and
    Mirror's even-size quirk included), kernel2 within the convolution tolerance, and both bit-identical to the Python host
    mirror, which does the same arithmetic with the same device convolutions.  Views of different, partly even extents."""
    from spim_registration_b200 import synthetic
    from spim_registration_b200.deconvolution import prepareKernels
    psfs = synthetic.make_psfs(3, (9, 7, 11), extracted=True)
    psfs[1] = np.ascontiguousarray(psfs[1][:, :6, :8])           # another extent, even in x and y
    k1, k2 = prepareKernels(psfs, psftype, 8)
    r1, r2 = o.init_kernels(psfs, psftype, 8)
    inp = _views(m, [np.ones((4, 4, 4), np.float32)] * 3, [1.0] * 3, psfs)
    inp.init(psftype, 8)
    for v, view in enumerate(inp.getViews()):
        np.testing.assert_array_equal(k1[v], r1[v])
        assert rel_l2(k2[v], r2[v]) <= 5e-6
        np.testing.assert_array_equal(k1[v], view.getKernel1())
        np.testing.assert_array_equal(k2[v], view.getKernel2())
    single = prepareKernels(psfs[:1], psftype, 8)                 # one view: kernel2 is the mirrored kernel1 for every type
    np.testing.assert_array_equal(single[1][0], o.compute_inverted_kernel(single[0][0]))


def test_convolution_cpu_symbol_runs_on_the_device(m, o, lib):
    """CUDASeparableConvolution.convolutionCPU (declared at CUDASeparableConvolution.java:19-21, called nowhere): kernels given
    by RADIUS, different per axis, one axis left alone -- computed by the device kernels behind convolve_N."""
    import spim_registration_b200._lib as L
    rng = np.random.default_rng(31)
    f = rng.random((30, 40, 50), dtype=np.float32) + np.float32(0.1)
    kx = rng.random(2 * 3 + 1).astype(np.float32)
    kz = rng.random(2 * 7 + 1).astype(np.float32)
    kx /= kx.sum()
    kz /= kz.sum()
    for oob, mode in ((0, "zero"), (1, "const"), (2, "clamp")):
        got = f.copy()
        lib.convolutionCPU(got.ctypes.data_as(L.c_float_p), kx.ctypes.data_as(L.c_float_p), None, kz.ctypes.data_as(L.c_float_p), 3, 0, 7,
                           50, 40, 30, oob, 0.6)
        assert rel_l2(got, o.separable_convolve(f, kx, None, kz, mode, 0.6)) <= CONV_TOL
    big = np.full(2 * 40 + 1, 1.0 / 81, np.float32)               # radius 40: beyond the direct kernels, FFT engine axis by axis
    got = f.copy()
    lib.convolutionCPU(got.ctypes.data_as(L.c_float_p), None, big.ctypes.data_as(L.c_float_p), None, 0, 40, 0, 50, 40, 30, 2, 0.0)
    assert rel_l2(got, o.separable_convolve(f, None, big, None, "clamp")) <= CONV_TOL


def test_peer_flag_barrier_on_one_gpu(m, o):
    """The neighbour-only flag barrier of the multi-process path (mvdExportFlags / mvdAttachPeerFlags / mvdPeerSignal /
    mvdPeerWait) on one GPU: every brick stores the step number into its neighbours' flag arrays, then every brick waits
    for its neighbours' -- signals before waits, because kernels of several bricks on one GPU that wait for one another may
    never run at the same time.  Same result as the copy exchange, bit for bit; a wait nobody signals for expires."""
    from spim_registration_b200 import synthetic
    from spim_registration_b200.parallel import LocalBrickGroup
    shape, V, grid = (37, 50, 150), 2, (2, 2, 2)
    ds = _dataset(o, shape, V, (9, 7, 11), seed=5)
    weights = synthetic.blending_weights(shape, V)
    k1, k2 = o.init_kernels(ds["psfs"], "EFFICIENT_BAYESIAN", 8)
    out = []
    for flags in (False, True):
        grp = LocalBrickGroup(shape[::-1], grid, ds["imgs"], weights, k1, k2, 0.006, use_flags=flags)
        try:
            grp.init_psi_average()
            if flags:
                grp.attach_peers()
            else:
                grp._exchange(0)
            grp.run_iterations(3)
            if flags:
                assert grp._barrier_count == 2 * 3 * V          # two exchanges per view and iteration went through the flags
            out.append(grp.get_psi())
        finally:
            grp.close()
    np.testing.assert_array_equal(out[0], out[1])


# ---- the quotient epilogue divides like IEEE float division, special values included ----------------------------------------
@pytest.mark.parametrize("scale", [1.0, 1e-25, 1e25])
def test_quotient_epilogue_is_ieee_division_with_special_values(m, lib, scale):
    """The quotient epilogue of the 128-wide windows evaluates the division for a float4 quad without branches when all
    four voxels lie inside the fast path's range and falls back to the generic division otherwise (quotient_quad,
    device_common.cuh).  Both must give img / blurred correctly rounded, 1 where img <= 0 or NaN
    (MVDeconvolution.java:496-499).  blurred comes from the same transform through the STORE epilogue; scale moves it out
    of the fast range (every quad takes the fallback)."""
    import ctypes as C
    torch = pytest.importorskip("torch")
    from spim_registration_b200 import _lib as L
    from spim_registration_b200.parallel import ShardedDeconvolution
    shape = (40, 150, 260)                                   # two 128-windows in x and y
    rng = np.random.default_rng(5)
    psi = ((rng.random(shape, dtype=np.float32) + np.float32(0.5)) * np.float32(scale)).astype(np.float32)
    k = _kernel((5, 9, 9), 3)
    img = (rng.random(shape, dtype=np.float32) * np.float32(100.0)).astype(np.float32)
    flat = img.reshape(-1)
    special = np.array([0.0, -1.0, 1e-30, 1e30, np.inf, np.nan, 1e-42, -0.0, 3e38, 2.0 ** -60, 2.0 ** 60], np.float32)
    pos = rng.choice(flat.size, size=special.size * 200, replace=False)
    flat[pos] = np.tile(special, 200)
    dec = ShardedDeconvolution(shape[::-1], [img], [1.0], [k], [k], 0.0, None, 0)
    try:
        L.check(lib.mvdSetDirectPath(dec.ctx, 0), "mvdSetDirectPath")
        dec.set_psi(psi)
        L.check(lib.mvdStepQuotient(dec.ctx, 0), "mvdStepQuotient")
        torch.cuda.synchronize()
        q = dec._arrays[1][0].cpu().numpy()
    finally:
        dec.close()
    blurred = m.convolve(psi, k, "mirror")                   # the same kernels up to the epilogue
    with np.errstate(all="ignore"):
        want = np.where(img > 0, img / blurred, np.float32(1.0)).astype(np.float32)
    same = (q == want) | (np.isnan(q) & np.isnan(want))
    assert same.mean() >= 0.999, "quotient differs from IEEE division in %.4f of the voxels" % (1 - same.mean())
    bad = ~same
    if bad.any():                                            # a last-bit difference of the two transforms, not of the division
        with np.errstate(all="ignore"):
            assert np.all(np.abs(q[bad] - want[bad]) <= 2e-7 * np.abs(want[bad]))
    assert (q[~(img > 0)] == 1).all()
    assert np.isinf(q[np.isinf(img)]).all()
