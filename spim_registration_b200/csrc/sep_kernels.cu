// Direct separable 3-D convolution for small rank-1 kernels (sm_100a).
//
// When a view's kernel is separable (K[z][y][x] = a[z] b[y] c[x], e.g. Gaussian PSFs and the compound
// kernels built from them) and at most 31 wide, the convolution does not need a Fourier transform:
//
//   sep_xy_kernel   x and y passes fused.  A CTA owns 128 columns and marches down a chunk of rows:
//                   every row (out-of-bounds rule applied on the fly) is staged once in shared
//                   memory, each thread forms the x convolution of its column from it and keeps the
//                   last KMAX x-convolved rows in a register ring for the y convolution.
//   sep_z_kernel    z pass + the same fused epilogues as the FFT path (store / quotient / update
//                   with Tikhonov, weights, statistics): a thread marches along z with a register
//                   ring of KMAX planes.
//
// Traffic per output voxel and convolution: read source (x halo overlap), write + read one float
// intermediate, epilogue operands -- about 22 B against about 50 B for the FFT path, and about
// 7 K instructions per voxel instead of 250.  The same kernels serve the reference's second native
// hook (in-place separable Gaussian convolution, CUDASeparableConvolution.java:13-17).
//
// Convention (same as the FFT path): out(p) = sum_j k[j] f(p - j + K/2) per axis.
// The host passes per axis KMAX "window taps": w[i] multiplies f(p - KMAX/2 + i).
#include "sep_kernels.h"
#include "device_common.cuh"

#include <algorithm>

namespace mvd {

namespace {

constexpr int SEP_NT = 128;   // threads per CTA = columns per CTA
constexpr int SEP_CHY = 64;   // rows per CTA in the xy kernel
constexpr int SEP_CHZ = 128;  // planes per CTA in the z kernel

template <int KMAX>
__global__ void __launch_bounds__(SEP_NT) sep_xy_kernel(Volume src, SepTaps taps, float* __restrict__ tmp, int bx, int by) {
    // tmp: [src.dim[2]][by][bx] -- every resident plane of the source (brick and ghost planes)
    constexpr int H = KMAX / 2;
    __shared__ float row[SEP_NT + KMAX - 1];
    __shared__ int xmap[SEP_NT + KMAX - 1];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * SEP_NT, y0 = blockIdx.y * SEP_CHY, za = blockIdx.z;
    const float fillv = src.ext == EXT_CONST ? src.ext_value : 0.f;
    for (int i = tid; i < SEP_NT + KMAX - 1; i += SEP_NT)
        xmap[i] = map_coord(x0 - H + i, src.dim[0], src.off[0], src.gmin[0], src.gdim[0], src.ext);
    float sx = 0.f;  // x convolution of a constant row = fill * sum of the x taps
#pragma unroll
    for (int i = 0; i < KMAX; ++i) sx += taps.x[i];
    const float* __restrict__ zbase = src.data + (size_t)za * src.dim[1] * (size_t)src.dim[0];
    float ring[KMAX];
#pragma unroll
    for (int i = 0; i < KMAX; ++i) ring[i] = 0.f;
    const int yend = min(y0 + SEP_CHY, by);
    __syncthreads();
#pragma unroll 1
    for (int r = y0 - H; r < yend + H; ++r) {
        const int yl = map_coord(r, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
        float xc;
        if (yl < 0) {
            xc = yl == -1 ? (taps.per_pass_fill ? fillv : fillv * sx) : 0.f;
        } else {
            const float* __restrict__ rp = zbase + (size_t)yl * src.dim[0];
            for (int i = tid; i < SEP_NT + KMAX - 1; i += SEP_NT) {
                const int xi = xmap[i];
                row[i] = xi >= 0 ? __ldg(rp + xi) : (xi == -1 ? fillv : 0.f);
            }
            __syncthreads();
            xc = 0.f;
#pragma unroll
            for (int i = 0; i < KMAX; ++i) xc = fmaf(taps.x[i], row[tid + i], xc);
            __syncthreads();  // the row buffer is rewritten by the next iteration
        }
#pragma unroll
        for (int i = 0; i + 1 < KMAX; ++i) ring[i] = ring[i + 1];
        ring[KMAX - 1] = xc;
        const int y = r - H;  // the ring now holds rows y - H .. y + H
        if (y >= y0 && x0 + tid < bx) {
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < KMAX; ++i) acc = fmaf(taps.y[i], ring[i], acc);
            tmp[((size_t)za * by + y) * bx + x0 + tid] = acc;
        }
    }
}

template <int KMAX, int EPI>
__global__ void __launch_bounds__(SEP_NT) sep_z_kernel(const float* __restrict__ tmp, Volume src, SepTaps taps, EpilogueArgs ep, int bx,
                                                       int by, int bz) {
    constexpr int H = KMAX / 2;
    __shared__ double s_sum[SEP_NT / 32];
    __shared__ float s_max[SEP_NT / 32];
    const int tid = threadIdx.x;
    const int x = blockIdx.x * SEP_NT + tid, y = blockIdx.y, z0 = blockIdx.z * SEP_CHZ;
    const bool live = x < bx;
    const float fillv = src.ext == EXT_CONST ? src.ext_value : 0.f;
    float sxy = 0.f, sy = 0.f;  // a plane of constant fill convolved in x and y
#pragma unroll
    for (int i = 0; i < KMAX; ++i) { sxy += taps.x[i]; sy += taps.y[i]; }
    const float fill_t = taps.per_pass_fill ? fillv : fillv * sxy * sy;
    const bool use_tik = ep.lambda > 0.0;
    const double two_lambda = 2.0 * ep.lambda, inv_lambda = use_tik ? 1.0 / ep.lambda : 0.0;
    double lsum = 0.0;
    float lmax = 0.f;
    float ring[KMAX];
#pragma unroll
    for (int i = 0; i < KMAX; ++i) ring[i] = 0.f;
    const int zend = min(z0 + SEP_CHZ, bz);
    const size_t col = (size_t)y * bx + (live ? x : 0);
    const size_t plane = (size_t)by * bx;
#pragma unroll 4
    for (int r = z0 - H; r < zend + H; ++r) {
        const int zl = map_coord(r, src.dim[2], src.off[2], src.gmin[2], src.gdim[2], src.ext);
        const float v = zl >= 0 ? __ldg(tmp + (size_t)zl * plane + col) : (zl == -1 ? fill_t : 0.f);
#pragma unroll
        for (int i = 0; i + 1 < KMAX; ++i) ring[i] = ring[i + 1];
        ring[KMAX - 1] = v;
        const int z = r - H;
        if (z >= z0 && live) {
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < KMAX; ++i) acc = fmaf(taps.z[i], ring[i], acc);
            acc += ep.add_const;
            const size_t idx = ((size_t)(z + ep.out_off[2]) * ep.out_dim[1] + (y + ep.out_off[1])) * (size_t)ep.out_dim[0] + x + ep.out_off[0];
            const size_t aidx = ((size_t)z * ep.aux_dim[1] + y) * (size_t)ep.aux_dim[0] + x;
            if constexpr (EPI == EPI_STORE) {
                ep.out[idx] = acc;
            } else if constexpr (EPI == EPI_QUOTIENT) {
                ep.out[idx] = quotient_value(__ldg(ep.img + aidx), acc);
            } else {
                const float last = ep.out[idx];
                const float w = ep.weight ? __ldg(ep.weight + aidx) : ep.weight_const;
                const float nv = next_psi_value(last, acc, w, use_tik, two_lambda, inv_lambda, ep.min_value);
                ep.out[idx] = nv;
                const float ch = fabsf(__fsub_rn(nv, last));
                lsum += (double)ch;
                lmax = fmaxf(lmax, ch);
            }
        }
    }
    if constexpr (EPI == EPI_UPDATE) {
        if (ep.sum_change) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
                lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
            }
            if ((tid & 31) == 0) { s_sum[tid >> 5] = lsum; s_max[tid >> 5] = lmax; }
            __syncthreads();
            if (tid == 0) {
                double s = 0.0;
                float mx = 0.f;
                for (int i = 0; i < SEP_NT / 32; ++i) { s += s_sum[i]; mx = fmaxf(mx, s_max[i]); }
                atomicAdd(ep.sum_change, s);
                atomicMax(ep.max_change_bits, __float_as_uint(mx));
            }
        }
    }
}

}  // namespace

// Window taps of a 1-D kernel of length K (<= kmax): w[i] multiplies f(p - kmax/2 + i), i.e. the
// coefficient of offset o = i - kmax/2 is k[K/2 - o].
void make_window_taps(const float* k, int K, int kmax, float* w) {
    const int H = kmax / 2;
    for (int i = 0; i < kmax; ++i) {
        const int j = K / 2 - (i - H);
        w[i] = (j >= 0 && j < K) ? k[j] : 0.f;
    }
}

int sep_kmax_for(int kx, int ky, int kz) {
    const int k = std::max(kx, std::max(ky, kz));
    return k <= 7 ? 7 : (k <= 15 ? 15 : (k <= 31 ? 31 : 0));
}

size_t sep_tmp_elems(const Volume& src, const int bdim[3]) { return (size_t)src.dim[2] * bdim[1] * (size_t)bdim[0]; }

cudaError_t launch_separable(const Volume& src, const int bdim[3], const SepTaps& taps, int kmax, int epilogue, const EpilogueArgs& ep,
                             float* tmp, cudaStream_t st) {
    const int bx = bdim[0], by = bdim[1], bz = bdim[2];
    const dim3 g1((bx + SEP_NT - 1) / SEP_NT, (by + SEP_CHY - 1) / SEP_CHY, src.dim[2]);
    const dim3 g2((bx + SEP_NT - 1) / SEP_NT, by, (bz + SEP_CHZ - 1) / SEP_CHZ);
    if (g1.y > 65535u || g1.z > 65535u || g2.y > 65535u || g2.z > 65535u) return cudaErrorInvalidValue;
#define SEP_LAUNCH(KM)                                                                                                  \
    {                                                                                                                   \
        sep_xy_kernel<KM><<<g1, SEP_NT, 0, st>>>(src, taps, tmp, bx, by);                                               \
        if (epilogue == EPI_STORE) sep_z_kernel<KM, EPI_STORE><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz);    \
        else if (epilogue == EPI_QUOTIENT) sep_z_kernel<KM, EPI_QUOTIENT><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz); \
        else sep_z_kernel<KM, EPI_UPDATE><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz);                        \
    }
    if (kmax == 7) SEP_LAUNCH(7)
    else if (kmax == 15) SEP_LAUNCH(15)
    else if (kmax == 31) SEP_LAUNCH(31)
    else return cudaErrorInvalidValue;
#undef SEP_LAUNCH
    return cudaGetLastError();
}

}  // namespace mvd
