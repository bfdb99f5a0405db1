// Direct separable 3-D convolution (see sep_kernels.cu).
#pragma once
#include "conv_types.h"

#include <cstddef>

namespace mvd {

constexpr int kSepMaxTaps = 31;

// per-axis window taps, zero-padded to the kernel's KMAX (7, 15 or 31): w[i] multiplies f(p - KMAX/2 + i)
struct SepTaps {
    float x[kSepMaxTaps], y[kSepMaxTaps], z[kSepMaxTaps];
    // 0: out = conv3d(ext(src)) -- a constant outside value is itself convolved by the earlier passes (the
    //    deconvolution loop); 1: every pass sees the plain outside value (three in-place 1-D passes, the
    //    separable-convolution hook)
    int per_pass_fill;
};

void make_window_taps(const float* k, int K, int kmax, float* w);
int sep_kmax_for(int kx, int ky, int kz);  // 7 / 15 / 31, or 0 when a kernel is too long for the direct path
size_t sep_tmp_elems(const Volume& src, const int bdim[3]);
// out = epilogue(conv(ext(src), x (x) y (x) z)) on the brick [0, bdim); tmp holds sep_tmp_elems floats
cudaError_t launch_separable(const Volume& src, const int bdim[3], const SepTaps& taps, int kmax, int epilogue, const EpilogueArgs& ep,
                             float* tmp, cudaStream_t st);

}  // namespace mvd
