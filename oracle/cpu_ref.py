"""ctypes face of oracle/libcpuref.so (cpu_ref.cpp) -- TEST INFRASTRUCTURE ONLY.

The C++ restatement follows the reference's own FFT-convolution route in float32 (pad to a
prime-factor-friendly size, rotate the kernel, R2C/C2C, multiply with or without the complex
conjugate, inverse, crop: J/fftMethods/FFTConvolution.java:470-551); oracle_np.py evaluates the
convolution FORMULA in float64.  The two are written independently and tests/test_oracle.py
holds them against each other.  numpy arrays are float32 [z, y, x].
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libcpuref.so")
_lib = None
F32P = C.POINTER(C.c_float)
EXT = {"zero": 0, "const": 1, "mirror": 2}


def load(build=True):
    global _lib
    if _lib is not None:
        return _lib
    src = os.path.join(_HERE, "cpu_ref.cpp")
    if build and (not os.path.exists(_PATH) or os.path.getmtime(_PATH) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    lib = C.CDLL(_PATH)
    lib.ref_fft_convolve.restype = C.c_int
    lib.ref_fft_convolve.argtypes = [F32P, C.POINTER(C.c_int), F32P, C.POINTER(C.c_int), C.c_int, C.c_float, C.c_int, F32P, C.POINTER(C.c_int)]
    lib.ref_run_iterations.restype = C.c_int
    lib.ref_nfft_small_complex.restype = C.c_int
    lib.ref_nfft_small_real.restype = C.c_int
    lib.ref_num_threads.restype = C.c_int
    lib.ref_compute_final_values.argtypes = [F32P, F32P, F32P, C.c_float, C.c_double, C.c_float, C.c_longlong, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.ref_compute_quotient.argtypes = [F32P, F32P, C.c_longlong]
    _lib = lib
    return lib


def _i3(shape_zyx):
    return (C.c_int * 3)(int(shape_zyx[2]), int(shape_zyx[1]), int(shape_zyx[0]))


def fft_convolve(f, kernel, mode, value=0.0, conjugate=False):
    """FFTConvolution.convolve on the extended accessor `mode` of f; returns (out, padded_xyz)."""
    lib = load()
    f = np.ascontiguousarray(f, dtype=np.float32)
    k = np.ascontiguousarray(kernel, dtype=np.float32)
    out = np.empty_like(f)
    pad = (C.c_int * 3)()
    rc = lib.ref_fft_convolve(f.ctypes.data_as(F32P), _i3(f.shape), k.ctypes.data_as(F32P), _i3(k.shape), EXT[mode], float(value),
                              int(bool(conjugate)), out.ctypes.data_as(F32P), pad)
    if rc != 0:
        raise RuntimeError("ref_fft_convolve failed")
    return out, tuple(pad)


def run_deconvolution(imgs, weights, kernel1, kernel2, num_iterations, lam, psi0, min_value=1e-4, conjugate=False):
    """MVDeconvolution.runIteration x num_iterations on the whole image (CPU branch).  Returns (psi, stats[it][v] = (sum, max))."""
    lib = load()
    V = len(imgs)
    imgs = [np.ascontiguousarray(i, dtype=np.float32) for i in imgs]
    k1 = [np.ascontiguousarray(k, dtype=np.float32) for k in kernel1]
    k2 = [np.ascontiguousarray(k, dtype=np.float32) for k in kernel2]
    wa = [None if np.isscalar(w) else np.ascontiguousarray(w, dtype=np.float32) for w in weights]
    wc = (C.c_float * V)(*[float(w) if np.isscalar(w) else 1.0 for w in weights])
    P = F32P * V
    null = C.cast(None, F32P)
    psi = np.array(psi0, dtype=np.float32, copy=True)
    kd1 = (C.c_int * (3 * V))(*[int(x) for k in k1 for x in k.shape[::-1]])
    kd2 = (C.c_int * (3 * V))(*[int(x) for k in k2 for x in k.shape[::-1]])
    stats = (C.c_double * (2 * V * num_iterations))()
    rc = lib.ref_run_iterations(P(*[i.ctypes.data_as(F32P) for i in imgs]), P(*[null if w is None else w.ctypes.data_as(F32P) for w in wa]), wc,
                                P(*[k.ctypes.data_as(F32P) for k in k1]), P(*[k.ctypes.data_as(F32P) for k in k2]), kd1, kd2, V, _i3(psi.shape),
                                int(num_iterations), C.c_double(lam), C.c_float(min_value), int(bool(conjugate)), psi.ctypes.data_as(F32P), stats)
    if rc != 0:
        raise RuntimeError("ref_run_iterations failed")
    st = np.array(stats[:]).reshape(num_iterations, V, 2)
    return psi, [[(st[i, v, 0], st[i, v, 1]) for v in range(V)] for i in range(num_iterations)]
