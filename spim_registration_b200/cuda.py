"""Host mirror of the reference package spim.process.cuda for the hot path.

Block / BlockGeneratorFixedSizePrecise reproduce the reference's tiling bit-exactly through the
native library (mvdDivideIntoBlocks); CUDAFourierConvolution is the proxy of the JNA interface of
the same name; CUDATools lists devices the way the reference's dialog does.  Arrays are numpy
float32 indexed [z, y, x]; size vectors are (x, y, z) like the reference's long[].
"""
import ctypes as C

import numpy as np

from . import _lib


class Block:
    """J/spim/process/cuda/Block.java:27-91 (fields) -- plain holder."""

    __slots__ = ("blockSize", "offset", "effectiveSize", "effectiveOffset", "effectiveLocalOffset", "isPrecise")

    def __init__(self, blockSize, offset, effectiveSize, effectiveOffset, effectiveLocalOffset, isPrecise=True):
        self.blockSize = list(blockSize)
        self.offset = list(offset)
        self.effectiveSize = list(effectiveSize)
        self.effectiveOffset = list(effectiveOffset)
        self.effectiveLocalOffset = list(effectiveLocalOffset)
        self.isPrecise = isPrecise

    def getBlockSize(self):
        return list(self.blockSize)

    def as_dict(self):
        return {k: getattr(self, k) for k in self.__slots__[:5]}

    # Block.copyBlock (Block.java:114-153): block[p] = extendedSource(p + offset).  `source` is a
    # callable (z0,z1,y0,y1,x0,x1) -> array returning the extended source on that window, or an
    # already extended array together with its origin.
    def copyBlock(self, extended, origin):
        o = [self.offset[d] - origin[d] for d in range(3)]
        b = self.blockSize
        return np.ascontiguousarray(extended[o[2]:o[2] + b[2], o[1]:o[1] + b[1], o[0]:o[0] + b[0]], dtype=np.float32)

    # Block.pasteBlock (Block.java:155-195)
    def pasteBlock(self, target, block):
        eo, es, el = self.effectiveOffset, self.effectiveSize, self.effectiveLocalOffset
        target[eo[2]:eo[2] + es[2], eo[1]:eo[1] + es[1], eo[0]:eo[0] + es[0]] = \
            block[el[2]:el[2] + es[2], el[1]:el[1] + es[1], el[0]:el[0] + es[0]]


class BlockGeneratorFixedSizePrecise:
    """J/spim/process/cuda/BlockGeneratorFixedSizePrecise.java:13-101."""

    def __init__(self, blockSize):
        self.blockSize = [int(b) for b in blockSize]

    def divideIntoBlocks(self, imgSize, kernelSize):
        lib = _lib.load()
        n = len(imgSize)
        LL = C.c_longlong * n
        img, ker, blk = LL(*map(int, imgSize)), LL(*map(int, kernelSize)), LL(*self.blockSize)
        total = lib.mvdDivideIntoBlocks(img, ker, blk, n, None, 0)
        if total == 0:
            return None  # the reference returns null (:38-42)
        if total < 0:
            raise _lib.MVDeconError("mvdDivideIntoBlocks failed")
        out = (C.c_longlong * (total * 5 * n))()
        got = lib.mvdDivideIntoBlocks(img, ker, blk, n, out, total)
        assert got == total
        a = np.frombuffer(out, dtype=np.int64).reshape(total, 5, n)
        return [Block(*[r.tolist() for r in rec]) for rec in a]


class CUDADevice:
    """J/spim/process/cuda/CUDADevice.java:3-38."""

    def __init__(self, deviceId, deviceName, totalDeviceMemory, freeDeviceMemory, major, minor):
        self.deviceId, self.deviceName = deviceId, deviceName
        self.totalDeviceMemory, self.freeDeviceMemory = totalDeviceMemory, freeDeviceMemory
        self.computeCapabilityMajor, self.computeCapabilityMinor = major, minor

    def __repr__(self):
        return "CUDADevice(%d, %r, %d MB, sm_%d%d)" % (self.deviceId, self.deviceName, self.totalDeviceMemory // (1 << 20),
                                                        self.computeCapabilityMajor, self.computeCapabilityMinor)


class CUDATools:
    """Device enumeration as in J/spim/process/cuda/CUDATools.java:37-98."""

    @staticmethod
    def queryCUDADetails():
        lib = _lib.load()
        n = lib.getNumDevicesCUDA()
        if n == -1:
            raise _lib.MVDeconError("Querying CUDA devices crashed, no devices available.")  # CUDATools.java:39-43
        devs = []
        for i in range(n):
            name = C.create_string_buffer(256)  # CUDATools.java:55
            lib.getNameDeviceCUDA(i, name)
            devs.append(CUDADevice(i, name.value.decode(), lib.getMemDeviceCUDA(i), lib.getFreeMemDeviceCUDA(i),
                                   lib.getCUDAcomputeCapabilityMajorVersion(i), lib.getCUDAcomputeCapabilityMinorVersion(i)))
        return devs


class CUDAFourierConvolution:
    """Proxy of the JNA interface J/spim/process/cuda/CUDAFourierConvolution.java:4-11."""

    def __init__(self, path=None):
        self._lib = _lib.load(path)

    def convolution3DfftCUDAInPlace(self, im, imDim, kernel, kernelDim, devCUDA):
        """im, kernel: contiguous float32 arrays (modified in place); imDim/kernelDim = (z,y,x)
        exactly as MVDeconFFTThreads.getCUDACoordinates passes them (MVDeconFFTThreads.java:136-144)."""
        assert im.dtype == np.float32 and im.flags.c_contiguous and kernel.dtype == np.float32 and kernel.flags.c_contiguous
        d = (C.c_int * 3)(*map(int, imDim))
        k = (C.c_int * 3)(*map(int, kernelDim))
        self._lib.convolution3DfftCUDAInPlace(im.ctypes.data_as(_lib.c_float_p), d, kernel.ctypes.data_as(_lib.c_float_p), k, int(devCUDA))
        code, msg = _lib.last_error()
        if code:
            raise _lib.MVDeconError("convolution3DfftCUDAInPlace: " + msg)


class CUDASeparableConvolution:
    """Proxy of the JNA interface J/spim/process/cuda/CUDASeparableConvolution.java:9-22 (the
    reference's second native hook: Gaussian blurs of the Difference-of-Gaussian detection)."""

    SIZES = (7, 15, 31, 63, 127)  # CUDASeparableConvolutionFunctions.java:14

    def __init__(self, path=None):
        self._lib = _lib.load(path)

    def convolve(self, size, image, kernelX, kernelY, kernelZ, imageW, imageH, imageD, convolveX, convolveY, convolveZ,
                 outofbounds, outofboundsvalue, devCUDA):
        """convolve_<size>: in-place; image float32 [d,h,w] contiguous; outofbounds 0 zero, 1 value, 2 extend last pixel.
        Returns the native boolean."""
        assert size in self.SIZES and image.dtype == np.float32 and image.flags.c_contiguous

        def ptr(k, on):
            if not on:
                return None
            assert k is not None and k.dtype == np.float32 and k.size == size
            return k.ctypes.data_as(_lib.c_float_p)
        fn = getattr(self._lib, "convolve_%d" % size)
        ok = fn(image.ctypes.data_as(_lib.c_float_p), ptr(kernelX, convolveX), ptr(kernelY, convolveY), ptr(kernelZ, convolveZ),
                int(imageW), int(imageH), int(imageD), int(bool(convolveX)), int(bool(convolveY)), int(bool(convolveZ)),
                int(outofbounds), float(outofboundsvalue), int(devCUDA))
        if not ok:
            _lib.last_error()  # read (and clear) the latch; the reference only sees the boolean
        return bool(ok)

    def convolve_7(self, *a):
        return self.convolve(7, *a)

    def convolve_15(self, *a):
        return self.convolve(15, *a)

    def convolve_31(self, *a):
        return self.convolve(31, *a)

    def convolve_63(self, *a):
        return self.convolve(63, *a)

    def convolve_127(self, *a):
        return self.convolve(127, *a)


class CUDASeparableConvolutionFunctions:
    """Host side of the hook, J/spim/process/cuda/CUDASeparableConvolutionFunctions.java."""

    ZERO, VALUE, EXTEND_BORDER_PIXELS = 0, 1, 2  # enum OutOfBounds (:13), passed as its ordinal (:178)

    @staticmethod
    def createGaussianKernel1DDouble(sigma, normalize=True):
        """mpicbg.imglib.util.Util.createGaussianKernel1DDouble (imglib1, a dependency outside the
        reference tree): size = max(3, 2 * round(3 sigma + 0.5) + 1), exp(-x^2 / (2 sigma^2)), sum 1."""
        if sigma <= 0:
            return np.ones(3, np.float64) * np.array([0.0, 1.0, 0.0])
        size = max(3, 2 * int(3 * sigma + 0.5) + 1)
        x = np.arange(size, dtype=np.float64) - size // 2
        k = np.exp(-(x * x) / (2.0 * sigma * sigma))
        return k / k.sum() if normalize else k

    @classmethod
    def getCUDAKernels(cls, sigma, supported=CUDASeparableConvolution.SIZES):
        """:199-229 -- one padded float kernel per dimension, all of the smallest supported size that holds them."""
        ks = [cls.createGaussianKernel1DDouble(s, True) for s in sigma]
        longest = max(len(k) for k in ks)
        sizes = [s for s in supported if longest <= s]
        if not sizes:
            return None
        size = min(sizes)
        out = []
        for k in ks:  # getFloatKernelPadded, :231-245
            p = np.zeros(size, np.float32)
            s0 = (size - len(k)) // 2
            p[s0:s0 + len(k)] = k.astype(np.float32)
            out.append(p)
        return out

    @classmethod
    def gauss(cls, img, sigma, oobs=2, oobsValue=0.0, cuda=None, cudaDeviceId=0):
        """:127-197 -- in-place Gaussian blur of img (float32 [z,y,x] / [y,x] / [x]); sigma per dimension in (x,y,z) order."""
        if img is None or sigma is None or len(sigma) != img.ndim or img.ndim > 3 or img.ndim == 0:
            return False
        if cudaDeviceId < 0:
            raise _lib.MVDeconError("gauss: CPU computation (device -1) is not provided, no CPU path")
        ks = cls.getCUDAKernels(sigma)
        if ks is None:
            return False
        cuda = cuda or CUDASeparableConvolution()
        dims = list(img.shape[::-1]) + [1] * (3 - img.ndim)
        k3 = ks + [None] * (3 - len(ks))
        return cuda.convolve(len(ks[0]), img, k3[0], k3[1], k3[2], dims[0], dims[1], dims[2], k3[0] is not None, k3[1] is not None,
                             k3[2] is not None, oobs, oobsValue, cudaDeviceId)


def convolve(src, kernel, ext_mode="zero", ext_value=0.0, device=0):
    """Resident-engine convolution of a whole volume (mvdConvolve)."""
    lib = _lib.load()
    src = np.ascontiguousarray(src, dtype=np.float32)
    kernel = np.ascontiguousarray(kernel, dtype=np.float32)
    out = np.empty_like(src)
    mode = {"zero": 0, "const": 1, "mirror": 2}[ext_mode]
    rc = lib.mvdConvolve(device, src.ctypes.data_as(_lib.c_float_p), _lib.int3(src.shape[::-1]),
                         kernel.ctypes.data_as(_lib.c_float_p), _lib.int3(kernel.shape[::-1]), mode, float(ext_value),
                         out.ctypes.data_as(_lib.c_float_p))
    _lib.check(rc, "mvdConvolve")
    return out


def plan_tiles(dim_xyz, kernel_xyz):
    lib = _lib.load()
    n = lib.mvdPlanTiles(_lib.int3(dim_xyz), _lib.int3(kernel_xyz), None, 0)
    if n < 0:
        raise _lib.MVDeconError("mvdPlanTiles: " + _lib.last_error()[1])
    buf = (C.c_int * (12 * n))()
    lib.mvdPlanTiles(_lib.int3(dim_xyz), _lib.int3(kernel_xyz), buf, n)
    return np.frombuffer(buf, dtype=np.int32).reshape(n, 4, 3).copy()
