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
