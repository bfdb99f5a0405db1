"""ctypes binding of libmvdecon.so (include/mvdecon.h).

This is the Python twin of the JNA proxies the reference uses
(J/spim/process/cuda/CUDAFourierConvolution.java:4-11, CUDAStandardFunctions.java:5-24,
loaded by NativeLibraryTools.java:62-131).  There is no fallback: if the shared library is
missing the import of anything that computes fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MVD_LIB_PATH") or os.path.join(_HERE, "libmvdecon.so")  # MVD_LIB_PATH: A/B builds during tuning

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)
c_double_p = C.POINTER(C.c_double)
c_ll_p = C.POINTER(C.c_longlong)

# name -> (restype, argtypes); every symbol declared in include/mvdecon.h
PROTOTYPES = {
    "convolve_7": (C.c_int, [c_float_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]),
    "convolve_15": (C.c_int, [c_float_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]),
    "convolve_31": (C.c_int, [c_float_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]),
    "convolve_63": (C.c_int, [c_float_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]),
    "convolve_127": (C.c_int, [c_float_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]),
    "getCUDAcomputeCapabilityMinorVersion": (C.c_int, [C.c_int]),
    "getCUDAcomputeCapabilityMajorVersion": (C.c_int, [C.c_int]),
    "getNumDevicesCUDA": (C.c_int, []),
    "getNameDeviceCUDA": (None, [C.c_int, C.c_char_p]),
    "getMemDeviceCUDA": (C.c_longlong, [C.c_int]),
    "getFreeMemDeviceCUDA": (C.c_longlong, [C.c_int]),
    "convolution3DfftCUDAInPlace": (None, [c_float_p, c_int_p, c_float_p, c_int_p, C.c_int]),
    "convolution3DfftCUDA": (C.c_void_p, [c_float_p, c_int_p, c_float_p, c_int_p, C.c_int]),
    "mvdGetLastError": (C.c_int, [C.c_char_p, C.c_int]),
    "mvdVersion": (C.c_char_p, []),
    "mvdCreate": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, c_int_p, c_int_p, c_int_p, c_int_p, C.c_double, C.c_float]),
    "mvdDestroy": (C.c_int, [C.c_void_p]),
    "mvdSetStream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mvdSynchronize": (C.c_int, [C.c_void_p]),
    "mvdAddView": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, c_float_p, c_int_p, c_float_p, c_int_p, C.c_int, c_int_p]),
    "mvdNumViews": (C.c_int, [C.c_void_p]),
    "mvdFirstIterationSums": (C.c_int, [C.c_void_p, c_double_p, c_ll_p]),
    "mvdFillPsi": (C.c_int, [C.c_void_p, C.c_float]),
    "mvdInitPsiAverage": (C.c_int, [C.c_void_p, c_double_p]),
    "mvdSetPsi": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "mvdGetPsi": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "mvdRunIterations": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p]),
    "mvdStepQuotient": (C.c_int, [C.c_void_p, C.c_int]),
    "mvdStepUpdate": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "mvdGetStats": (C.c_int, [C.c_void_p, C.c_int, C.c_int, c_double_p, c_double_p]),
    "mvdProfileStart": (C.c_int, [C.c_void_p]),
    "mvdProfileStop": (C.c_int, [C.c_void_p, c_double_p, c_ll_p, c_double_p]),
    "mvdArrayInfo": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), c_int_p, c_int_p]),
    "mvdTransformView": (C.c_int, [C.c_int, C.c_void_p, c_int_p, c_double_p, c_int_p, c_int_p, c_float_p, c_float_p, C.c_float, C.c_void_p,
                                   C.c_void_p, C.c_int, C.c_int]),
    "mvdNormalizeWeights": (C.c_int, [C.c_int, C.POINTER(C.c_void_p), C.c_int, C.c_longlong, C.c_double, c_int_p, c_double_p, C.c_int]),
    "mvdVirtualWeightSum": (C.c_int, [C.c_int, C.c_int, c_int_p, c_double_p, c_float_p, c_float_p, c_int_p, c_int_p, C.c_void_p, c_int_p, c_double_p,
                                      C.c_int]),
    "mvdVirtualWeight": (C.c_int, [C.c_int, c_int_p, c_double_p, c_float_p, c_float_p, c_int_p, c_int_p, C.c_void_p, C.c_double, C.c_void_p, C.c_int]),
    "mvdExtractPSF": (C.c_int, [C.c_int, C.c_void_p, c_int_p, c_double_p, C.c_int, c_int_p, C.c_int, c_float_p, C.c_int]),
    "mvdTransformedPSFSize": (C.c_int, [c_int_p, c_double_p, c_int_p]),
    "mvdTransformPSF": (C.c_int, [C.c_int, c_float_p, c_int_p, c_double_p, c_int_p, c_float_p]),
    "mvdSetDirectPath": (C.c_int, [C.c_void_p, C.c_int]),
    "mvdViewUsesDirectPath": (C.c_int, [C.c_void_p, C.c_int, c_int_p, c_int_p]),
    "mvdExportArrays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "mvdAttachPeer": (C.c_int, [C.c_void_p, c_int_p, c_int_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "mvdDetachPeers": (C.c_int, [C.c_void_p]),
    "mvdMaskNeverUpdated": (C.c_int, [C.c_void_p]),
    "mvdConvolve": (C.c_int, [C.c_int, c_float_p, c_int_p, c_float_p, c_int_p, C.c_int, C.c_float, c_float_p]),
    "mvdConvolveDevice": (C.c_int, [C.c_int, C.c_void_p, c_int_p, c_float_p, c_int_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "mvdPlanTiles": (C.c_int, [c_int_p, c_int_p, c_int_p, C.c_int]),
    "mvdDivideIntoBlocks": (C.c_longlong, [c_ll_p, c_ll_p, c_ll_p, C.c_int, c_ll_p, C.c_longlong]),
    "mvdKernelLaunchCount": (C.c_longlong, []),
    "mvdSetTuning": (C.c_int, [C.c_char_p, C.c_int]),
    "mvdExportFlags": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "mvdAttachPeerFlags": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "mvdPeerSignal": (C.c_int, [C.c_void_p, C.c_uint]),
    "mvdPeerWait": (C.c_int, [C.c_void_p, C.c_uint]),
    "mvdPeerBarrier": (C.c_int, [C.c_void_p, C.c_uint]),
    "mvdPrepareKernels": (C.c_int, [C.c_int, C.c_int, C.POINTER(c_float_p), c_int_p, C.c_int, C.c_int, C.POINTER(c_float_p), C.POINTER(c_float_p)]),
    "convolutionCPU": (None, [c_float_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]),
    "mvdSeparableConvolveNoAlloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, c_int_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int, C.c_float,
                                              C.c_void_p]),
    "mvdCreateMulti": (C.c_int, [C.POINTER(C.c_void_p), c_int_p, C.c_int, c_int_p, c_int_p, C.c_double, C.c_float]),
    "mvdDestroyMulti": (C.c_int, [C.c_void_p]),
    "mvdMultiNumBricks": (C.c_int, [C.c_void_p]),
    "mvdMultiBrickInfo": (C.c_int, [C.c_void_p, C.c_int, c_int_p, c_int_p, c_int_p]),
    "mvdMultiAddView": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, c_float_p, c_int_p, c_float_p, c_int_p]),
    "mvdMultiSetDirectPath": (C.c_int, [C.c_void_p, C.c_int]),
    "mvdMultiInitPsiAverage": (C.c_int, [C.c_void_p, c_double_p]),
    "mvdMultiSetPsi": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mvdMultiRunIterations": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p]),
    "mvdMultiMaskNeverUpdated": (C.c_int, [C.c_void_p]),
    "mvdMultiGetPsi": (C.c_int, [C.c_void_p, C.c_void_p]),
}

_lib = None


class MVDeconError(RuntimeError):
    pass


def load(path=None):
    """Load libmvdecon.so (like Native.loadLibrary, NativeLibraryTools.java:124).  Raises if absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise MVDeconError(
            "libmvdecon.so not found at %s -- build it with `make -C spim_registration_b200/csrc` "
            "(or __graft_entry__.build()); there is no CPU fallback" % p)
    lib = C.CDLL(p)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def last_error():
    buf = C.create_string_buffer(1024)
    code = load().mvdGetLastError(buf, 1024)
    return code, buf.value.decode("utf-8", "replace")


def check(rc, what):
    if rc != 0:
        code, msg = last_error()
        raise MVDeconError("%s failed (%d): %s" % (what, rc, msg))


def int3(v):
    return (C.c_int * 3)(int(v[0]), int(v[1]), int(v[2]))
