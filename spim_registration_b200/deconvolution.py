"""Host mirror of spim.process.fusion.deconvolution for the hot path, on top of libmvdecon.so.

Same class names, argument meaning and behaviour as the reference (MVDeconInput.java:10-59,
MVDeconFFT.java:25-535, MVDeconvolution.java:36-705); the arithmetic of the loop runs entirely
in the CUDA library through the resident C ABI (include/mvdecon.h part 2).  numpy arrays are
float32 indexed [z, y, x].
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from .cuda import convolve as _device_convolve

F32 = np.float32


class PSFTYPE:  # MVDeconFFT.java:28
    OPTIMIZATION_II = "OPTIMIZATION_II"
    OPTIMIZATION_I = "OPTIMIZATION_I"
    EFFICIENT_BAYESIAN = "EFFICIENT_BAYESIAN"
    INDEPENDENT = "INDEPENDENT"
    ALL = (OPTIMIZATION_II, OPTIMIZATION_I, EFFICIENT_BAYESIAN, INDEPENDENT)


# ---- host-side kernel arithmetic (runs on the Java side of the boundary in the reference) ----
def _portions(size, n):  # FusionHelper.divideIntoPortions, FusionHelper.java:236-260
    chunk, mod = divmod(size, n)
    return [(i * chunk, chunk + (mod if i == n - 1 else 0)) for i in range(n)]


class AdjustInput:
    """J/spim/process/fusion/deconvolution/AdjustInput.java."""

    @staticmethod
    def sumImg(img, numThreads):
        # AdjustInput.java:44-100; sums[0] is counted twice (:93-97) -- kept, it is what the
        # reference's kernels are normalised with.
        flat = np.asarray(img, dtype=F32).ravel().astype(np.float64)
        parts = [math.fsum(flat[s:s + n]) for s, n in _portions(flat.size, numThreads * 2)]
        return math.fsum([parts[0]] + parts)

    @staticmethod
    def normImg(img, numThreads):
        # AdjustInput.java:29-36 : t = (float)((double)t / sum)
        s = AdjustInput.sumImg(img, numThreads)
        return (np.asarray(img, dtype=F32).astype(np.float64) / s).astype(F32)


class Mirror:
    """J/spim/process/fusion/deconvolution/Mirror.java:31-108."""

    @staticmethod
    def mirror(image, dimension):
        a = np.array(image, dtype=F32, copy=True)
        ax = a.ndim - 1 - dimension
        n = a.shape[ax]
        idx = np.arange(n)[::-1].copy()
        if n % 2 == 0:  # positions dim/2-1 and dim/2 are swapped twice (Mirror.java:72)
            idx[n // 2 - 1], idx[n // 2] = n // 2 - 1, n // 2
        return np.ascontiguousarray(np.take(a, idx, axis=ax))


class MVDeconFFT:
    """One view: image, weight, PSF (MVDeconFFT.java:57-158).  `weight` may be an array or a
    scalar (the constant-weight case)."""

    def __init__(self, image, weight, kernel, deviceList=(0,), useBlocks=False, blockSize=None, saveMemory=False):
        self.image = np.ascontiguousarray(image, dtype=F32)
        if np.isscalar(weight):
            self.weight = float(weight)
        else:
            self.weight = np.ascontiguousarray(weight, dtype=F32)
            assert self.weight.shape == self.image.shape
        self.kernel1 = np.ascontiguousarray(kernel, dtype=F32)
        self.kernel2 = None
        self.deviceList = list(deviceList)
        self.device0 = self.deviceList[0]
        if self.device0 < 0:
            raise _lib.MVDeconError("device -1 (CPU) is not available in this library: there is no CPU path")
        self.useBlocks, self.blockSize, self.saveMemory = useBlocks, blockSize, saveMemory
        self.numViews = 0

    def setNumViews(self, n):
        self.numViews = n

    def getImage(self):
        return self.image

    def getWeight(self):
        return self.weight

    def getKernel1(self):
        return self.kernel1

    def getKernel2(self):
        return self.kernel2

    @staticmethod
    def computeInvertedKernel(kernel):  # MVDeconFFT.java:315-323
        k = np.asarray(kernel, dtype=F32)
        for d in range(k.ndim):
            k = Mirror.mirror(k, d)
        return k

    @staticmethod
    def computeExponentialKernel(kernel, numViews):  # MVDeconFFT.java:305-313,325-333
        k = np.asarray(kernel, dtype=F32)
        r = k.copy()
        for _ in range(1, numViews):
            r = (r * k).astype(F32)
        return r

    def _conv0(self, inp, kernel):
        # FFTConvolution(Views.extendZero(input), input, Views.extendZero(kernel), kernel, output)
        return _device_convolve(inp, kernel, "zero", 0.0, self.device0)

    def init(self, iterationType, views, numThreads=8):
        """MVDeconFFT.init (MVDeconFFT.java:162-303)."""
        self.kernel1 = AdjustInput.normImg(self.kernel1, numThreads)
        self.iterationType = iterationType
        if self.numViews == 0:
            self.numViews = 1
        inv = MVDeconFFT.computeInvertedKernel
        if self.numViews == 1 or iterationType == PSFTYPE.INDEPENDENT:
            self.kernel2 = inv(self.kernel1)
        elif iterationType == PSFTYPE.EFFICIENT_BAYESIAN:
            tmp = inv(self.kernel1)
            for view in views:
                if view is self:
                    continue
                out = self._conv0(inv(self.kernel1), view.kernel1)
                out = self._conv0(out, inv(view.kernel1))
                tmp = (out * tmp).astype(F32)
            self.kernel2 = AdjustInput.normImg(tmp, numThreads)
        elif iterationType == PSFTYPE.OPTIMIZATION_I:
            tmp = self.kernel1.copy()
            for view in views:
                if view is self:
                    continue
                out = self._conv0(self.kernel1, inv(view.kernel1))
                tmp = (out * tmp).astype(F32)
            tmp = AdjustInput.normImg(tmp, numThreads)
            self.kernel2 = inv(tmp)
        else:  # OPTIMIZATION_II
            e = MVDeconFFT.computeExponentialKernel(self.kernel1, self.numViews)
            e = AdjustInput.normImg(e, numThreads)
            self.kernel2 = inv(e)


def prepareKernels(psfs, iterationType, numThreads=8, device=0):
    """MVDeconInput.init through the library (mvdPrepareKernels): (kernel1[], kernel2[]) for the views' PSFs, for hosts that
    do not carry their own MVDeconFFT.init.  Same values as MVDeconInput.init below (bit for bit: the same host arithmetic,
    the same device convolutions)."""
    lib = _lib.load()
    psfs = [np.ascontiguousarray(p, dtype=F32) for p in psfs]
    V = len(psfs)
    k1 = [np.empty_like(p) for p in psfs]
    k2 = [np.empty_like(p) for p in psfs]
    P = _lib.c_float_p * V
    dims = (C.c_int * (3 * V))(*[int(x) for p in psfs for x in p.shape[::-1]])
    _lib.check(lib.mvdPrepareKernels(int(device), V, P(*[p.ctypes.data_as(_lib.c_float_p) for p in psfs]), dims, PSFTYPE.ALL.index(iterationType),
                                     int(numThreads), P(*[k.ctypes.data_as(_lib.c_float_p) for k in k1]),
                                     P(*[k.ctypes.data_as(_lib.c_float_p) for k in k2])), "mvdPrepareKernels")
    return k1, k2


class MVDeconInput:
    """MVDeconInput.java:10-59."""

    minValue = F32(0.0001)

    def __init__(self):
        self.views = []

    def add(self, view):
        self.views.append(view)
        for v in self.views:
            v.setNumViews(self.getNumViews())

    def init(self, iterationType, numThreads=8):
        for view in self.views:  # in order: later views are still un-normalised (MVDeconInput.java:43-49)
            view.init(iterationType, self.views, numThreads)
        return self

    def getViews(self):
        return self.views

    def getNumViews(self):
        return len(self.views)


class MVDeconvolution:
    """MVDeconvolution.java:73-190: the constructor runs the whole deconvolution; getPsi() returns
    the result.  `initialImage` plays the role of the static MVDeconvolution.initialImage."""

    minValue = F32(0.0001)  # MVDeconvolution.java:49

    def __init__(self, views, iterationType, numIterations, lam, osemspeedup=1.0, osemspeedupindex=0, name="",
                 initialImage=None, device=None, numThreads=8, prepared=False, directPath=True, debug=False, debugInterval=1):
        lib = _lib.load()
        self.views, self.name, self.lam = views, name, float(lam)
        self.data = views.getViews()
        self.numViews = len(self.data)
        if not prepared:
            views.init(iterationType, numThreads)
        dev = self.data[0].device0 if device is None else device
        shape = self.data[0].getImage().shape
        devs = list(self.data[0].deviceList) if device is None else [device]
        if len(devs) > 1:
            # several devices (MVDeconFFT.java:426-445: one thread per device): the volume is sharded over them behind
            # one handle of the C ABI, one process (include/mvdecon.h, mvdCreateMulti)
            self._run_multi(lib, devs, shape, numIterations, initialImage, directPath, debug, debugInterval)
            return
        self._ctx = C.c_void_p()
        _lib.check(lib.mvdCreate(C.byref(self._ctx), dev, _lib.int3(shape[::-1]), None, None, None, self.lam, float(self.minValue)),
                   "mvdCreate")
        try:
            for v in self.data:
                w = v.getWeight()
                wp, wc = (None, float(w)) if isinstance(w, float) else (w.ctypes.data_as(C.c_void_p), 1.0)
                k1, k2 = v.getKernel1(), v.getKernel2()
                _lib.check(lib.mvdAddView(self._ctx, v.getImage().ctypes.data_as(C.c_void_p), wp, wc,
                                          k1.ctypes.data_as(_lib.c_float_p), _lib.int3(k1.shape[::-1]),
                                          k2.ctypes.data_as(_lib.c_float_p), _lib.int3(k2.shape[::-1]), 0, None), "mvdAddView")
            # which convolutions take the direct separable path (kernel class 7 / 15 / 31, 0 = FFT engine)
            _lib.check(lib.mvdSetDirectPath(self._ctx, int(bool(directPath))), "mvdSetDirectPath")
            self.directPath = []
            for i in range(self.numViews):
                a, b = C.c_int(), C.c_int()
                _lib.check(lib.mvdViewUsesDirectPath(self._ctx, i, C.byref(a), C.byref(b)), "mvdViewUsesDirectPath")
                self.directPath.append((a.value, b.value))
            if initialImage is not None:
                init = np.ascontiguousarray(initialImage, dtype=F32)
                _lib.check(lib.mvdSetPsi(self._ctx, init.ctypes.data_as(C.c_void_p), 0), "mvdSetPsi")
                self.avg = None
            else:
                avg = C.c_double()
                _lib.check(lib.mvdInitPsiAverage(self._ctx, C.byref(avg)), "mvdInitPsiAverage")
                self.avg = avg.value
            n = numIterations * self.numViews
            s, m = (C.c_double * max(n, 1))(), (C.c_double * max(n, 1))()
            # debug mode (MVDeconvolution.java:139-175): a copy of psi every debugInterval iterations, before the final mask
            self.debugStack = []
            if debug:
                for it in range(numIterations):
                    s1, m1 = (C.c_double * self.numViews)(), (C.c_double * self.numViews)()
                    _lib.check(lib.mvdRunIterations(self._ctx, 1, s1, m1), "mvdRunIterations")
                    s[it * self.numViews:(it + 1) * self.numViews] = s1[:]
                    m[it * self.numViews:(it + 1) * self.numViews] = m1[:]
                    if it % max(1, debugInterval) == 0:
                        snap = np.empty(shape, dtype=F32)
                        _lib.check(lib.mvdGetPsi(self._ctx, snap.ctypes.data_as(C.c_void_p), 0), "mvdGetPsi")
                        self.debugStack.append((it, snap))
            else:
                _lib.check(lib.mvdRunIterations(self._ctx, numIterations, s, m), "mvdRunIterations")
            self.sumChange = np.array(s[:n]).reshape(numIterations, self.numViews)
            self.maxChange = np.array(m[:n]).reshape(numIterations, self.numViews)
            self.i = numIterations
            _lib.check(lib.mvdMaskNeverUpdated(self._ctx), "mvdMaskNeverUpdated")
            self.psi = np.empty(shape, dtype=F32)
            _lib.check(lib.mvdGetPsi(self._ctx, self.psi.ctypes.data_as(C.c_void_p), 0), "mvdGetPsi")
        finally:
            lib.mvdDestroy(self._ctx)
            self._ctx = None

    def _run_multi(self, lib, devs, shape, numIterations, initialImage, directPath, debug, debugInterval):
        kmax = [1, 1, 1]
        for v in self.data:
            for k in (v.getKernel1(), v.getKernel2()):
                kmax = [max(a, b) for a, b in zip(kmax, k.shape[::-1])]
        h = C.c_void_p()
        _lib.check(lib.mvdCreateMulti(C.byref(h), (C.c_int * len(devs))(*devs), len(devs), _lib.int3(shape[::-1]), _lib.int3(kmax), self.lam,
                                      float(self.minValue)), "mvdCreateMulti")
        try:
            for v in self.data:
                w = v.getWeight()
                wp, wc = (None, float(w)) if isinstance(w, float) else (w.ctypes.data_as(C.c_void_p), 1.0)
                k1, k2 = v.getKernel1(), v.getKernel2()
                _lib.check(lib.mvdMultiAddView(h, v.getImage().ctypes.data_as(C.c_void_p), wp, wc, k1.ctypes.data_as(_lib.c_float_p),
                                               _lib.int3(k1.shape[::-1]), k2.ctypes.data_as(_lib.c_float_p), _lib.int3(k2.shape[::-1])), "mvdMultiAddView")
            _lib.check(lib.mvdMultiSetDirectPath(h, int(bool(directPath))), "mvdMultiSetDirectPath")
            self.directPath = None
            self.bricks = []
            for b in range(lib.mvdMultiNumBricks(h)):
                d, off, dim = C.c_int(), (C.c_int * 3)(), (C.c_int * 3)()
                _lib.check(lib.mvdMultiBrickInfo(h, b, C.byref(d), off, dim), "mvdMultiBrickInfo")
                self.bricks.append((d.value, tuple(off), tuple(dim)))
            if initialImage is not None:
                init = np.ascontiguousarray(initialImage, dtype=F32)
                _lib.check(lib.mvdMultiSetPsi(h, init.ctypes.data_as(C.c_void_p)), "mvdMultiSetPsi")
                self.avg = None
            else:
                avg = C.c_double()
                _lib.check(lib.mvdMultiInitPsiAverage(h, C.byref(avg)), "mvdMultiInitPsiAverage")
                self.avg = avg.value
            nv = self.numViews
            n = numIterations * nv
            s, m = (C.c_double * max(n, 1))(), (C.c_double * max(n, 1))()
            self.debugStack = []
            if debug:
                for it in range(numIterations):
                    s1, m1 = (C.c_double * nv)(), (C.c_double * nv)()
                    _lib.check(lib.mvdMultiRunIterations(h, 1, s1, m1), "mvdMultiRunIterations")
                    s[it * nv:(it + 1) * nv] = s1[:]
                    m[it * nv:(it + 1) * nv] = m1[:]
                    if it % max(1, debugInterval) == 0:
                        snap = np.empty(shape, dtype=F32)
                        _lib.check(lib.mvdMultiGetPsi(h, snap.ctypes.data_as(C.c_void_p)), "mvdMultiGetPsi")
                        self.debugStack.append((it, snap))
            else:
                _lib.check(lib.mvdMultiRunIterations(h, numIterations, s, m), "mvdMultiRunIterations")
            self.sumChange = np.array(s[:n]).reshape(numIterations, nv)
            self.maxChange = np.array(m[:n]).reshape(numIterations, nv)
            self.i = numIterations
            _lib.check(lib.mvdMultiMaskNeverUpdated(h), "mvdMultiMaskNeverUpdated")
            self.psi = np.empty(shape, dtype=F32)
            _lib.check(lib.mvdMultiGetPsi(h, self.psi.ctypes.data_as(C.c_void_p)), "mvdMultiGetPsi")
        finally:
            lib.mvdDestroyMulti(h)

    def getPsi(self):
        return self.psi

    def getCurrentIteration(self):
        return self.i
