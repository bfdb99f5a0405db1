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
