// Inputs of the deconvolution loop, produced on the device: the step the
// reference runs on the CPU immediately before the loop.
//
//   transform_view_kernel   TransformInputAndWeights.loop (TransformInputAndWeights.java:81-114):
//                           for every voxel of the output bounding box map the world coordinate
//                           into the view (inverse affine), sample the view with n-linear
//                           interpolation over the mirror-single extended image
//                           (Views.interpolate(Views.extendMirrorSingle(img), NLinear), :62), keep
//                           max(minValue, value) inside the view and 0 outside (0 = no data,
//                           MVDeconvolution.java:496), and evaluate the cosine blending weight of
//                           BlendingRealRandomAccess.computeWeight (:71-101) at the same position.
//   normalize_weights_kernel WeightNormalizer.ApplyDirectly (WeightNormalizer.java:133-181): every
//                           weight divided by the sum over the views (0 / 0 stays NaN exactly as in
//                           the reference), plus the overlap statistics (min and average number of
//                           views with weight > 0).
//   osem_kernel             ProcessForDeconvolution.adjustForOSEM (:351-362): w = min(1, w * speedup).
//
// One thread per output voxel, x fastest: the gathers of neighbouring threads fall into the same
// source rows, writes are coalesced.  The arithmetic that decides rounding is done in double like
// the Java code (AffineTransform3D.applyInverse, the interpolation weights, the weight sum).
#include "input_kernels.h"

#include <algorithm>
#include <cmath>

namespace mvd {

namespace {

__constant__ double c_blend_lut[1001];  // BlendingRealRandomAccess.lookUp (:25-36)

__device__ __forceinline__ int mirror_single(int i, int n) {
    // Views.extendMirrorSingle: ... 2 1 | 0 1 2 ... n-1 | n-2 n-3 ...
    if (n == 1) return 0;
    const int p = 2 * n - 2;
    i %= p;
    if (i < 0) i += p;
    return i < n ? i : p - i;
}

__device__ __forceinline__ float blend_weight(const float t[3], const int dim_minus1[3], const float border[3], const float blending[3]) {
    float min_distance = 1.f;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float l = t[d];  // the interval starts at 0
        const float dist = fmaxf(0.f, fminf(__fsub_rn(l, border[d]), __fsub_rn(__fsub_rn((float)dim_minus1[d], l), border[d])));
        if (dist == 0.f) return 0.f;
        const float rel = __fdiv_rn(dist, blending[d]);
        if (rel < 1.f) {
            const long long idx = (long long)floor((double)rel * 1000.0 + 0.5);  // Math.round
            min_distance = (float)((double)min_distance * c_blend_lut[idx]);       // float *= double
        }
    }
    return min_distance;
}

__global__ void __launch_bounds__(256) transform_view_kernel(const float* __restrict__ src, TransformArgs a, float* __restrict__ out_img,
                                                             float* __restrict__ out_w) {
    // grid: x = chunks of a row, y = rows (y + Y * z) in a grid-stride loop
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const long long rows = (long long)a.out_dim[1] * a.out_dim[2];
    if (x >= a.out_dim[0]) return;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int y = (int)(r % a.out_dim[1]), z = (int)(r / a.out_dim[1]);
        const size_t i = (size_t)r * a.out_dim[0] + x;
        // s = voxel + offset as floats, t = inverse * s evaluated in double and stored as float (applyInverse(float[], float[]))
        const double s0 = (double)(float)(x + a.offset[0]), s1 = (double)(float)(y + a.offset[1]), s2 = (double)(float)(z + a.offset[2]);
        float t[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) t[d] = (float)(a.inv[4 * d] * s0 + a.inv[4 * d + 1] * s1 + a.inv[4 * d + 2] * s2 + a.inv[4 * d + 3]);
        float v = 0.f;
        if (t[0] >= 0.f && t[1] >= 0.f && t[2] >= 0.f && t[0] < (float)a.src_dim[0] && t[1] < (float)a.src_dim[1] && t[2] < (float)a.src_dim[2]) {
            int i0[3], i1[3];
            double w[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                const double f = floor((double)t[d]);
                w[d] = (double)t[d] - f;
                i0[d] = mirror_single((int)f, a.src_dim[d]);
                i1[d] = mirror_single((int)f + 1, a.src_dim[d]);
            }
            const size_t sx = a.src_dim[0], sxy = (size_t)a.src_dim[0] * a.src_dim[1];
            auto at = [&](int xx, int yy, int zz) { return (double)__ldg(src + xx + sx * yy + sxy * zz); };
            const double c00 = at(i0[0], i0[1], i0[2]) * (1.0 - w[0]) + at(i1[0], i0[1], i0[2]) * w[0];
            const double c10 = at(i0[0], i1[1], i0[2]) * (1.0 - w[0]) + at(i1[0], i1[1], i0[2]) * w[0];
            const double c01 = at(i0[0], i0[1], i1[2]) * (1.0 - w[0]) + at(i1[0], i0[1], i1[2]) * w[0];
            const double c11 = at(i0[0], i1[1], i1[2]) * (1.0 - w[0]) + at(i1[0], i1[1], i1[2]) * w[0];
            const double c0 = c00 * (1.0 - w[1]) + c10 * w[1];
            const double c1 = c01 * (1.0 - w[1]) + c11 * w[1];
            v = fmaxf(a.min_value, (float)(c0 * (1.0 - w[2]) + c1 * w[2]));
        }
        out_img[i] = v;
        if (out_w) {
            const int dm1[3] = {a.src_dim[0] - 1, a.src_dim[1] - 1, a.src_dim[2] - 1};
            out_w[i] = blend_weight(t, dm1, a.border, a.blending);
        }
    }
}

__global__ void __launch_bounds__(256) normalize_weights_kernel(WeightPtrs w, size_t n, unsigned long long* __restrict__ d_count,
                                                                int* __restrict__ d_min) {
    __shared__ unsigned long long s_cnt[8];
    __shared__ int s_min[8];
    unsigned long long lcnt = 0;
    int lmin = w.n;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        double sum = 0.0;
        int count = 0;
        for (int v = 0; v < w.n; ++v) {
            const float x = w.w[v][i];
            sum += (double)x;
            if (x > 0.f) ++count;
        }
        lcnt += count;
        lmin = min(lmin, count);
        for (int v = 0; v < w.n; ++v) w.w[v][i] = (float)((double)w.w[v][i] / sum);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lcnt += __shfl_xor_sync(0xffffffffu, lcnt, o);
        lmin = min(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
    }
    if ((threadIdx.x & 31) == 0) { s_cnt[threadIdx.x >> 5] = lcnt; s_min[threadIdx.x >> 5] = lmin; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long c = 0;
        int m = w.n;
        for (int i = 0; i < 8; ++i) { c += s_cnt[i]; m = min(m, s_min[i]); }
        atomicAdd(d_count, c);
        atomicMin(d_min, m);
    }
}

// ---- virtual weights (the dialog's default weight type) -------------------------------------------
// The reference evaluates, per voxel, per view and per iteration, the blending function at the voxel's
// position inside the view (TransformedInterpolatedRealRandomAccess.applyInverse :68-95: integer world
// coordinate -> double, t0*i00 + t1*i01 + t2*i02 + i03 left to right, cast to float) divided by the
// float image of weight sums (NormalizingRandomAccess.get :36-45).  Here both are evaluated once and kept.
__device__ __forceinline__ void view_position(const double* __restrict__ inv, int wx, int wy, int wz, float t[3]) {
    const double t0 = (double)wx, t1 = (double)wy, t2 = (double)wz;
#pragma unroll
    for (int d = 0; d < 3; ++d)  // no fused multiply-add: Java rounds every product and sum
        t[d] = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(t0, inv[4 * d]), __dmul_rn(t1, inv[4 * d + 1])), __dmul_rn(t2, inv[4 * d + 2])), inv[4 * d + 3]);
}

// WeightNormalizer.ComputeSumImage (:183-232): double sum of the views' float weights, stored as float
// where it exceeds 1 and as 1 elsewhere; number of views with weight > 0 per voxel (min and total).
__global__ void __launch_bounds__(256) virtual_weight_sum_kernel(const BlendView* __restrict__ views, int n_views, VolumeBox box,
                                                                 float* __restrict__ sum_w, unsigned long long* __restrict__ d_count,
                                                                 int* __restrict__ d_min) {
    __shared__ unsigned long long s_cnt[8];
    __shared__ int s_min[8];
    unsigned long long lcnt = 0;
    int lmin = n_views;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const long long rows = (long long)box.dim[1] * box.dim[2];
    if (x < box.dim[0]) {
        for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
            const int y = (int)(r % box.dim[1]), z = (int)(r / box.dim[1]);
            double sum = 0.0;
            int count = 0;
            for (int v = 0; v < n_views; ++v) {
                const BlendView& bv = views[v];
                float t[3];
                view_position(bv.inv, x + box.offset[0], y + box.offset[1], z + box.offset[2], t);
                const float w = blend_weight(t, bv.dim_minus1, bv.border, bv.blending);
                sum += (double)w;
                if (w > 0.f) ++count;
            }
            lcnt += count;
            lmin = min(lmin, count);
            sum_w[(size_t)r * box.dim[0] + x] = sum > 1.0 ? (float)sum : 1.f;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lcnt += __shfl_xor_sync(0xffffffffu, lcnt, o);
        lmin = min(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
    }
    if ((threadIdx.x & 31) == 0) { s_cnt[threadIdx.x >> 5] = lcnt; s_min[threadIdx.x >> 5] = lmin; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long c = 0;
        int m = n_views;
        for (int i = 0; i < 8; ++i) { c += s_cnt[i]; m = min(m, s_min[i]); }
        atomicAdd(d_count, c);
        atomicMin(d_min, m);
    }
}

// NormalizingRandomAccess.get (:36-45): (float) min(1, (double)w / (double)sumW * osemspeedup)
__global__ void __launch_bounds__(256) virtual_weight_kernel(BlendView bv, VolumeBox box, const float* __restrict__ sum_w, double osem,
                                                             float* __restrict__ out_w) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const long long rows = (long long)box.dim[1] * box.dim[2];
    if (x >= box.dim[0]) return;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int y = (int)(r % box.dim[1]), z = (int)(r / box.dim[1]);
        const size_t i = (size_t)r * box.dim[0] + x;
        float t[3];
        view_position(bv.inv, x + box.offset[0], y + box.offset[1], z + box.offset[2], t);
        const float w = blend_weight(t, bv.dim_minus1, bv.border, bv.blending);
        const double v = __dmul_rn(__ddiv_rn((double)w, (double)__ldg(sum_w + i)), osem);
        out_w[i] = v != v ? (float)v : (float)fmin(1.0, v);  // Math.min(1, NaN) is NaN
    }
}

__global__ void __launch_bounds__(256) osem_kernel(float* __restrict__ w, size_t n, float speedup) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float v = __fmul_rn(w[i], speedup);
        w[i] = v != v ? v : fminf(1.f, v);  // Math.min(1, NaN) is NaN
    }
}

// ---- PSF extraction (ExtractPSF.java) -------------------------------------------------------------
// n-linear sample of `img` at a double position; mode 0: zero outside (Views.extendZero), 1: periodic
__device__ __forceinline__ double sample_linear(const float* __restrict__ img, const int dim[3], double px, double py, double pz, int mode) {
    const double p[3] = {px, py, pz};
    int i0[3], i1[3];
    double w[3];
    bool in0[3], in1[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const double f = floor(p[d]);
        w[d] = p[d] - f;
        const long long a = (long long)f, b = a + 1;
        if (mode == 1) {
            long long ma = a % dim[d], mb = b % dim[d];
            if (ma < 0) ma += dim[d];
            if (mb < 0) mb += dim[d];
            i0[d] = (int)ma; i1[d] = (int)mb;
            in0[d] = in1[d] = true;
        } else {
            in0[d] = a >= 0 && a < dim[d];
            in1[d] = b >= 0 && b < dim[d];
            i0[d] = in0[d] ? (int)a : 0;
            i1[d] = in1[d] ? (int)b : 0;
        }
    }
    const size_t sx = dim[0], sxy = (size_t)dim[0] * dim[1];
    auto at = [&](bool okx, int xx, bool oky, int yy, bool okz, int zz) {
        return (okx && oky && okz) ? (double)__ldg(img + xx + sx * yy + sxy * zz) : 0.0;
    };
    const double c00 = at(in0[0], i0[0], in0[1], i0[1], in0[2], i0[2]) * (1.0 - w[0]) + at(in1[0], i1[0], in0[1], i0[1], in0[2], i0[2]) * w[0];
    const double c10 = at(in0[0], i0[0], in1[1], i1[1], in0[2], i0[2]) * (1.0 - w[0]) + at(in1[0], i1[0], in1[1], i1[1], in0[2], i0[2]) * w[0];
    const double c01 = at(in0[0], i0[0], in0[1], i0[1], in1[2], i1[2]) * (1.0 - w[0]) + at(in1[0], i1[0], in0[1], i0[1], in1[2], i1[2]) * w[0];
    const double c11 = at(in0[0], i0[0], in1[1], i1[1], in1[2], i1[2]) * (1.0 - w[0]) + at(in1[0], i1[0], in1[1], i1[1], in1[2], i1[2]) * w[0];
    const double c0 = c00 * (1.0 - w[1]) + c10 * w[1];
    const double c1 = c01 * (1.0 - w[1]) + c11 * w[1];
    return c0 * (1.0 - w[2]) + c1 * w[2];
}

// ExtractPSF.extractPSFLocal (:353-394): psf(v) = sum over the bead locations of the view sampled at
// v - size/2 + location (periodic extension), accumulated in float in the order of the locations
__global__ void __launch_bounds__(128) extract_psf_kernel(const float* __restrict__ img, PsfArgs a, const double* __restrict__ loc, int n_loc,
                                                          float* __restrict__ psf) {
    const int n = a.size[0] * a.size[1] * a.size[2];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int x = i % a.size[0], y = (i / a.size[0]) % a.size[1], z = i / (a.size[0] * a.size[1]);
    float acc = 0.f;
    for (int l = 0; l < n_loc; ++l) {
        const double px = (double)(x - a.size[0] / 2) + loc[3 * l], py = (double)(y - a.size[1] / 2) + loc[3 * l + 1],
                     pz = (double)(z - a.size[2] / 2) + loc[3 * l + 2];
        acc = __fadd_rn(acc, (float)sample_linear(img, a.dim, px, py, pz, 1));
    }
    psf[i] = acc;
}

// ExtractPSF.transform (:396-437): out(v) = psf sampled (zero outside) at inverse(v + offset)
__global__ void __launch_bounds__(128) transform_psf_kernel(const float* __restrict__ psf, PsfArgs a, float* __restrict__ out) {
    const int n = a.size[0] * a.size[1] * a.size[2];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int x = i % a.size[0], y = (i / a.size[0]) % a.size[1], z = i / (a.size[0] * a.size[1]);
    const double t0 = (double)x + a.offset[0], t1 = (double)y + a.offset[1], t2 = (double)z + a.offset[2];
    double p[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) p[d] = a.inv[4 * d] * t0 + a.inv[4 * d + 1] * t1 + a.inv[4 * d + 2] * t2 + a.inv[4 * d + 3];
    out[i] = (float)sample_linear(psf, a.dim, p[0], p[1], p[2], 0);
}

int grid_for(size_t n) {
    const size_t want = (n + 255) / 256;
    return (int)(want < (size_t)(148 * 16) ? (want ? want : 1) : (size_t)(148 * 16));
}

}  // namespace

cudaError_t upload_blend_table() {
    // lookUp[indexFor(d)] = (cos((1 - d) * pi) + 1) / 2 for d = 0, 0.001, ... (accumulated) while d <= 1.0001
    static double lut[1001];
    for (int i = 0; i <= 1000; ++i) lut[i] = 0.0;
    const double pi = 3.14159265358979323846;
    for (double d = 0; d <= 1.0001; d = d + 0.001) {
        const long long idx = (long long)std::floor(d * 1000.0 + 0.5);
        if (idx >= 0 && idx <= 1000) lut[idx] = (std::cos((1 - d) * pi) + 1) / 2;
    }
    return cudaMemcpyToSymbol(c_blend_lut, lut, sizeof(lut));
}

cudaError_t launch_transform_view(const float* src, const TransformArgs& a, float* out_img, float* out_w, cudaStream_t st) {
    const size_t n = (size_t)a.out_dim[0] * a.out_dim[1] * a.out_dim[2];
    if (n == 0) return cudaSuccess;
    const long long rows = (long long)a.out_dim[1] * a.out_dim[2];
    const dim3 grid((a.out_dim[0] + 255) / 256, (unsigned)(rows < 65535 ? rows : 65535));
    transform_view_kernel<<<grid, 256, 0, st>>>(src, a, out_img, out_w);
    return cudaGetLastError();
}

cudaError_t launch_normalize_weights(const WeightPtrs& w, size_t n, unsigned long long* d_count, int* d_min, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    normalize_weights_kernel<<<grid_for(n), 256, 0, st>>>(w, n, d_count, d_min);
    return cudaGetLastError();
}

cudaError_t launch_extract_psf(const float* img, const PsfArgs& a, const double* d_loc, int n_loc, float* d_psf, cudaStream_t st) {
    const int n = a.size[0] * a.size[1] * a.size[2];
    if (n <= 0) return cudaSuccess;
    extract_psf_kernel<<<(n + 127) / 128, 128, 0, st>>>(img, a, d_loc, n_loc, d_psf);
    return cudaGetLastError();
}

cudaError_t launch_transform_psf(const float* d_psf, const PsfArgs& a, float* d_out, cudaStream_t st) {
    const int n = a.size[0] * a.size[1] * a.size[2];
    if (n <= 0) return cudaSuccess;
    transform_psf_kernel<<<(n + 127) / 128, 128, 0, st>>>(d_psf, a, d_out);
    return cudaGetLastError();
}

static dim3 row_grid(const int dim[3]) {
    const long long rows = (long long)dim[1] * dim[2];
    return dim3((unsigned)((dim[0] + 255) / 256), (unsigned)std::min<long long>(rows, 65535), 1);
}

cudaError_t launch_virtual_weight_sum(const BlendView* d_views, int n_views, const VolumeBox& box, float* sum_w, unsigned long long* d_count,
                                      int* d_min, cudaStream_t st) {
    virtual_weight_sum_kernel<<<row_grid(box.dim), 256, 0, st>>>(d_views, n_views, box, sum_w, d_count, d_min);
    return cudaGetLastError();
}

cudaError_t launch_virtual_weight(const BlendView& view, const VolumeBox& box, const float* sum_w, double osem, float* out_w, cudaStream_t st) {
    virtual_weight_kernel<<<row_grid(box.dim), 256, 0, st>>>(view, box, sum_w, osem, out_w);
    return cudaGetLastError();
}

cudaError_t launch_osem(float* w, size_t n, float speedup, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    osem_kernel<<<grid_for(n), 256, 0, st>>>(w, n, speedup);
    return cudaGetLastError();
}

}  // namespace mvd
