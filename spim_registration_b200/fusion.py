"""Host side of the step before the loop: the views are resampled into the
bounding box of the deconvolved volume and get their blending weights on the device.

Mirrors J/spim/process/fusion/deconvolution/TransformInputAndWeights.java, TransformInput.java,
WeightNormalizer.java and ProcessForDeconvolution.adjustForOSEM; the arithmetic runs in
libmvdecon.so (csrc/input_kernels.cu), there is no CPU path.
"""
import ctypes as C

import numpy as np

from . import _lib

F32 = np.float32


def _inverse_3x4(transform):
    """AffineTransform3D.inverse(): transform is the view's registration (view -> world), 3x4 or 4x4."""
    m = np.eye(4, dtype=np.float64)
    t = np.asarray(transform, dtype=np.float64)
    m[:3, :4] = t[:3, :4]
    return np.ascontiguousarray(np.linalg.inv(m)[:3, :4])


def transformInputAndWeights(img, transform, offset, dimensions, blendingBorder=None, blendingRange=None, minValue=1e-4, device=0):
    """TransformInputAndWeights (with blending parameters) or TransformInput (without):
    img float32 [z,y,x]; transform: the view's registration (view -> world); offset / dimensions: the
    bounding box (x,y,z).  Returns (transformedImg, weightImg or None) as float32 [z,y,x]."""
    lib = _lib.load()
    img = np.ascontiguousarray(img, dtype=F32)
    inv = _inverse_3x4(transform)
    out = np.empty(tuple(dimensions[::-1]), F32)
    w = np.empty_like(out) if blendingBorder is not None else None
    fb = (C.c_float * 3)(*map(float, blendingBorder)) if blendingBorder is not None else None
    fr = (C.c_float * 3)(*map(float, blendingRange)) if blendingBorder is not None else None
    _lib.check(lib.mvdTransformView(int(device), img.ctypes.data_as(C.c_void_p), _lib.int3(img.shape[::-1]),
                                    inv.ctypes.data_as(_lib.c_double_p), _lib.int3(offset), _lib.int3(dimensions), fb, fr, float(minValue),
                                    out.ctypes.data_as(C.c_void_p), None if w is None else w.ctypes.data_as(C.c_void_p), 0, 0),
               "mvdTransformView")
    return out, w


class WeightNormalizer:
    """WeightNormalizer.java:52-131 with directly applied normalisation (PRECOMPUTED_WEIGHTS,
    ProcessForDeconvolution.java:304-307), optionally followed by adjustForOSEM (:351-362)."""

    def __init__(self, weights, device=0):
        self.weights = [np.ascontiguousarray(w, dtype=F32) for w in weights]
        self.device = device
        self.minOverlappingViews, self.avgOverlappingViews = None, None

    def process(self, osemspeedup=1.0):
        lib = _lib.load()
        ptrs = (C.c_void_p * len(self.weights))(*[w.ctypes.data_as(C.c_void_p).value for w in self.weights])
        mn, avg = C.c_int(), C.c_double()
        _lib.check(lib.mvdNormalizeWeights(int(self.device), ptrs, len(self.weights), int(self.weights[0].size), float(osemspeedup),
                                           C.byref(mn), C.byref(avg), 0), "mvdNormalizeWeights")
        self.minOverlappingViews, self.avgOverlappingViews = mn.value, avg.value
        return True

    def getMinOverlappingViews(self):
        return self.minOverlappingViews

    def getAvgOverlappingViews(self):
        return self.avgOverlappingViews


class VirtualWeights:
    """WeightType.VIRTUAL_WEIGHTS (ProcessForDeconvolution.java:204-208, 308-309, 375-379): blending functions
    seen through the views' registrations, normalised by a sum image (WeightNormalizer.java:42-46, 183-232,
    NormalizingRandomAccess.java:36-45).  The reference evaluates them on every access; here every view's
    weight image is evaluated once on the device (csrc/input_kernels.cu) with identical values.

    views: list of (srcDim (x,y,z), transform (view -> world, 3x4 or 4x4), blendingBorder, blendingRange)."""

    def __init__(self, views, offset, dimensions, device=0):
        self.views = [(tuple(int(v) for v in d), _inverse_3x4(t), tuple(map(float, b)), tuple(map(float, r))) for d, t, b, r in views]
        self.offset, self.dimensions, self.device = tuple(offset), tuple(dimensions), device
        self.sumWeights = None
        self.minOverlappingViews, self.avgOverlappingViews = None, None
        self.osemspeedup = 1.0

    def process(self):
        """WeightNormalizer.process with a sum image: computes sumWeights and the overlap statistics."""
        lib = _lib.load()
        n = len(self.views)
        dims = (C.c_int * (3 * n))(*[v for d, _, _, _ in self.views for v in d])
        inv = np.ascontiguousarray(np.stack([i for _, i, _, _ in self.views]).reshape(-1))
        fb = (C.c_float * (3 * n))(*[v for _, _, b, _ in self.views for v in b])
        fr = (C.c_float * (3 * n))(*[v for _, _, _, r in self.views for v in r])
        self.sumWeights = np.empty(self.dimensions[::-1], F32)
        mn, avg = C.c_int(), C.c_double()
        _lib.check(lib.mvdVirtualWeightSum(int(self.device), n, dims, inv.ctypes.data_as(_lib.c_double_p), fb, fr, _lib.int3(self.offset),
                                           _lib.int3(self.dimensions), self.sumWeights.ctypes.data_as(C.c_void_p), C.byref(mn), C.byref(avg), 0),
                   "mvdVirtualWeightSum")
        self.minOverlappingViews, self.avgOverlappingViews = mn.value, avg.value
        return True

    def setOSEMspeedup(self, osemspeedup):
        self.osemspeedup = float(osemspeedup)

    def weight(self, view):
        """The values NormalizingRandomAccess.get returns for one view, as float32 [z,y,x]."""
        lib = _lib.load()
        d, inv, b, r = self.views[view]
        out = np.empty(self.dimensions[::-1], F32)
        _lib.check(lib.mvdVirtualWeight(int(self.device), _lib.int3(d), inv.ctypes.data_as(_lib.c_double_p), (C.c_float * 3)(*b), (C.c_float * 3)(*r),
                                        _lib.int3(self.offset), _lib.int3(self.dimensions), self.sumWeights.ctypes.data_as(C.c_void_p),
                                        self.osemspeedup, out.ctypes.data_as(C.c_void_p), 0), "mvdVirtualWeight")
        return out


class ExtractPSF:
    """ExtractPSF.java: extraction of a view's PSF from bead locations and its transformation into the
    fused coordinate system; the sampling runs in libmvdecon.so (csrc/input_kernels.cu)."""

    def __init__(self, device=0):
        self.device = device
        self.pointSpreadFunctions, self.originalPSFs = [], []

    @staticmethod
    def _model(model):
        return np.ascontiguousarray(np.asarray(model, dtype=np.float64)[:3, :4])

    def extractPSFLocal(self, img, locations, size, normalize=False):
        """:353-394 (+ normalize, :277-295).  img [z,y,x]; locations [[x,y,z], ...]; size (x,y,z)."""
        lib = _lib.load()
        img = np.ascontiguousarray(img, dtype=F32)
        loc = np.ascontiguousarray(np.asarray(locations, dtype=np.float64).reshape(-1, 3))
        psf = np.empty(tuple(size[::-1]), F32)
        _lib.check(lib.mvdExtractPSF(int(self.device), img.ctypes.data_as(C.c_void_p), _lib.int3(img.shape[::-1]),
                                     loc.ctypes.data_as(_lib.c_double_p), len(loc), _lib.int3(size), int(bool(normalize)),
                                     psf.ctypes.data_as(_lib.c_float_p), 0), "mvdExtractPSF")
        return psf

    def transformPSF(self, psf, model):
        """:304-346.  model: the view's registration (view -> world), 3x4 or 4x4."""
        lib = _lib.load()
        psf = np.ascontiguousarray(psf, dtype=F32)
        m = self._model(model)
        new = (C.c_int * 3)()
        _lib.check(lib.mvdTransformedPSFSize(_lib.int3(psf.shape[::-1]), m.ctypes.data_as(_lib.c_double_p), new), "mvdTransformedPSFSize")
        out = np.empty((new[2], new[1], new[0]), F32)
        _lib.check(lib.mvdTransformPSF(int(self.device), psf.ctypes.data_as(_lib.c_float_p), _lib.int3(psf.shape[::-1]),
                                       m.ctypes.data_as(_lib.c_double_p), new, out.ctypes.data_as(_lib.c_float_p)), "mvdTransformPSF")
        return out

    def extractNextImg(self, img, model, locations, psfSize):
        """:256-275: extract, normalise, transform; keeps both PSFs."""
        original = self.extractPSFLocal(img, locations, psfSize, normalize=True)
        psf = self.transformPSF(original, model)
        self.originalPSFs.append(original)
        self.pointSpreadFunctions.append(psf)
        return psf

    def computeAverageTransformedPSF(self):
        """:173-215 (host arithmetic: a few thousand additions): the PSFs added point-mirrored about their centres."""
        psfs = self.pointSpreadFunctions
        size = [max(p.shape[2 - d] for p in psfs) for d in range(3)]
        avg = np.zeros(size[::-1], F32)
        ac = [s // 2 for s in size]
        for p in psfs:
            pz, py, px = p.shape
            flipped = p[::-1, ::-1, ::-1]  # loc = centre - loc + avgCentre
            # index of the flipped image's voxel 0: centre - (n - 1) + avgCentre
            o = [px // 2 - (px - 1) + ac[0], py // 2 - (py - 1) + ac[1], pz // 2 - (pz - 1) + ac[2]]
            sl_a, sl_p = [], []
            for d, n in zip(range(3), (px, py, pz)):
                lo, hi = max(0, o[d]), min(size[d], o[d] + n)
                sl_a.append(slice(lo, hi))
                sl_p.append(slice(lo - o[d], hi - o[d]))
            avg[sl_a[2], sl_a[1], sl_a[0]] += flipped[sl_p[2], sl_p[1], sl_p[0]]
        return avg
