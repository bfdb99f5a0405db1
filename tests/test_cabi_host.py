"""CPU tests of the C-ABI library and the host mirror: the shared library loads, exports every
symbol include/mvdecon.h declares, and its host-only entry points (block tiling, tile planner,
error latch) behave -- no compute call is made, no GPU is needed."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_np as o

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "mvdecon.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{]*\)\s*;", text)
    return sorted(set(names))


def test_header_and_binding_agree():
    from spim_registration_b200 import _lib
    declared = _declared_symbols()
    assert "convolution3DfftCUDAInPlace" in declared and "mvdRunIterations" in declared and len(declared) >= 30
    assert sorted(_lib.PROTOTYPES) == declared  # the ctypes twin binds exactly the header


def test_library_loads_and_exports_every_symbol(lib):
    for name in _declared_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.mvdVersion()


def test_legacy_symbol_names_match_the_jna_interfaces():
    """The JNA proxies bind by method name: the names below are the method names of
    CUDAStandardFunctions.java:15-23 and CUDAFourierConvolution.java:9-10."""
    declared = set(_declared_symbols())
    for n in ("getCUDAcomputeCapabilityMinorVersion", "getCUDAcomputeCapabilityMajorVersion", "getNumDevicesCUDA",
              "getNameDeviceCUDA", "getMemDeviceCUDA", "getFreeMemDeviceCUDA", "convolution3DfftCUDA", "convolution3DfftCUDAInPlace"):
        assert n in declared


def test_missing_library_fails_loudly(tmp_path):
    from spim_registration_b200 import _lib
    with pytest.raises(_lib.MVDeconError, match="no CPU fallback"):
        _lib.load(str(tmp_path / "libmvdecon.so"))


@pytest.mark.parametrize("img,ker,blk", [
    ([1024, 1024], [16, 32], [384, 384]),                   # Block.java:367-402
    ([1024, 1024, 512], [31, 31, 31], [256, 256, 256]),     # BASELINE config 3
    ([512, 512, 256], [31, 31, 31], [256, 256, 256]),
    ([130, 97, 41], [15, 11, 9], [64, 48, 40]),
    ([7, 5, 3], [3, 3, 3], [8, 8, 8]),                      # image smaller than one block
    ([100], [9], [32]),
])
def test_divide_into_blocks_bit_exact(img, ker, blk):
    """mvdDivideIntoBlocks == oracle restatement of BlockGeneratorFixedSizePrecise.java:25-101, every field."""
    import spim_registration_b200 as m
    got = m.BlockGeneratorFixedSizePrecise(blk).divideIntoBlocks(img, ker)
    want = o.divide_into_blocks(img, ker, blk)
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g.as_dict() == w


def test_divide_into_blocks_too_small_returns_none():
    import spim_registration_b200 as m
    assert m.BlockGeneratorFixedSizePrecise([30, 30, 30]).divideIntoBlocks([64, 64, 64], [31, 31, 31]) is None
    assert o.divide_into_blocks([64, 64, 64], [31, 31, 31], [30, 30, 30]) is None


def test_block_copy_paste_index_maps_bit_exact():
    """Block.copyBlock / pasteBlock of the host mirror == oracle copy/paste (pure index arithmetic)."""
    import spim_registration_b200 as m
    rng = np.random.default_rng(0)
    f = rng.random((20, 30, 40), dtype=np.float32)
    blocks = m.BlockGeneratorFixedSizePrecise([16, 16, 16]).divideIntoBlocks([40, 30, 20], [5, 7, 3])
    out_a, out_b = np.zeros_like(f), np.zeros_like(f)
    ext = o.extend(f, (16, 16, 16), (16, 16, 16), "mirror")
    for b in blocks:
        data = b.copyBlock(ext, (-16, -16, -16))
        np.testing.assert_array_equal(data, o.copy_block(f, b.as_dict(), "mirror"))
        b.pasteBlock(out_a, data)
        o.paste_block(out_b, data, b.as_dict())
    np.testing.assert_array_equal(out_a, f)
    np.testing.assert_array_equal(out_b, f)


@pytest.mark.parametrize("dim,k", [((1024, 1024, 512), (31, 31, 31)), ((128, 128, 128), (15, 15, 15)),
                                   ((70, 130, 150), (9, 11, 15)), ((5, 6, 7), (15, 15, 15)), ((33, 1, 41), (4, 1, 8))])
def test_tile_plan_covers_volume_exactly_once(dim, k):
    """Overlap-save plan (host-only): windows are supported FFT sizes, effective regions partition
    the volume, and every output's kernel support lies inside its window."""
    import spim_registration_b200 as m
    t = m.plan_tiles(dim, k)
    b, o_, e, n = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
    assert set(np.unique(b[:, :2])) <= {32, 64, 128} and set(np.unique(b[:, 2])) <= {16, 32, 64, 128, 256, 512}
    total = int(np.prod(n.astype(np.int64), axis=1).sum())
    assert total == dim[0] * dim[1] * dim[2]
    cover = np.zeros(dim[::-1], np.int8)
    for i in range(len(t)):
        cover[e[i, 2]:e[i, 2] + n[i, 2], e[i, 1]:e[i, 1] + n[i, 1], e[i, 0]:e[i, 0] + n[i, 0]] += 1
    assert (cover == 1).all()
    for d in range(3):
        hl, hr = k[d] - 1 - k[d] // 2, k[d] // 2
        assert (e[:, d] - hl >= o_[:, d]).all()                       # reads down to x - (K-1-K/2)
        assert (e[:, d] + n[:, d] - 1 + hr <= o_[:, d] + b[:, d] - 1).all()  # and up to x + K/2
        if d < 2:
            assert ((e[:, d] - o_[:, d]) % 4 == 0).all()              # float4-aligned effective regions


def test_kernel_too_large_for_a_window_is_an_error_not_a_crash():
    import spim_registration_b200 as m
    from spim_registration_b200 import _lib
    with pytest.raises(_lib.MVDeconError, match="kernel too large"):
        m.plan_tiles((64, 64, 64), (200, 3, 3))
    code, _ = _lib.last_error()
    assert code == 0  # reading the latch cleared it


def test_null_arguments_latch_errors(lib):
    from spim_registration_b200 import _lib
    assert lib.mvdRunIterations(None, 1, None, None) != 0
    code, msg = _lib.last_error()
    assert code != 0 and "null context" in msg
    assert lib.mvdNumViews(None) == -1
    assert lib.mvdDestroy(None) == 0


def test_host_kernel_arithmetic_matches_oracle():
    """AdjustInput / Mirror / exponent kernel of the host mirror (pure numpy, runs on the Java side
    of the boundary in the reference) == oracle restatement, bit for bit."""
    import spim_registration_b200 as m
    rng = np.random.default_rng(4)
    k = rng.random((6, 7, 9), dtype=np.float32)
    np.testing.assert_array_equal(m.AdjustInput.normImg(k, 8), o.norm_img(k, 8))
    assert m.AdjustInput.sumImg(k, 3) == o.sum_img(k, 3)
    for d in range(3):
        np.testing.assert_array_equal(m.Mirror.mirror(k, d), o.mirror(k, d))  # incl. the even-size quirk (z = 6)
    np.testing.assert_array_equal(m.MVDeconFFT.computeInvertedKernel(k), o.compute_inverted_kernel(k))
    np.testing.assert_array_equal(m.MVDeconFFT.computeExponentialKernel(k, 4), o.compute_exponential_kernel(k, 4))


def test_cpu_device_is_refused():
    import spim_registration_b200 as m
    from spim_registration_b200 import _lib
    with pytest.raises(_lib.MVDeconError, match="no CPU path"):
        m.MVDeconFFT(np.ones((4, 4, 4), np.float32), 1.0, np.ones((3, 3, 3), np.float32), deviceList=(-1,))
