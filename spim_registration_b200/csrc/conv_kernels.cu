// Tiled overlap-save 3-D FFT convolution for sm_100a: three kernels per tile batch.
//
//   fwd_xy : source volume (+ on-the-fly out-of-bounds rule) -> 2-D R2C FFT of one
//            z-plane of one tile in shared memory -> half-spectrum plane in scratch
//   z_pass : forward z FFT, pointwise multiply with the resident kernel spectrum,
//            inverse z FFT, in place on 16..256 plane positions x BZ per CTA
//   inv_xy : inverse 2-D C2R FFT of one effective z-plane in shared memory with the
//            fused per-voxel epilogue (store / quotient / multiplicative update)
//
// Replaces, for one block, copyBlock -> FFTConvolution/convolution3DfftCUDAInPlace ->
// pasteBlock of the reference (MVDeconFFTThreads.java:18-94, Block.java:114-359,
// J/fftMethods/FFTConvolution.java:470-551) and the elementwise passes
// MVDeconvolution.java:473-525 and :582-703.
#include "conv_kernels.h"
#include "fft_device.cuh"

namespace mvd {

static constexpr int NT = 256;  // threads per CTA of the xy kernels

// shared-memory plane: BY rows of H = BX/2 packed complex columns (+8 pad so the row
// stride is 8 mod 16 float2) and an XOR swizzle that makes both the Stockham scatter
// (stride-8 indices) and the unit-stride gathers conflict-free per half-warp.
template <int BX> struct PlaneLayout {
    static constexpr int H = BX / 2;
    static constexpr int SROW = H + 8;
    __device__ __forceinline__ static int at(int row, int i) { return row * SROW + (i ^ ((i >> 3) & 7)); }
};

// Map a local coordinate through the source's out-of-bounds rule.
// Returns the local array index, or -1 when the value is the fill value `fill`.
__device__ __forceinline__ int map_coord(int c, int dim, int off, int gmin, int gdim, int ext, float ext_value, float& fill) {
    int g = c + gmin;
    if (g < 0 || g >= gdim) {
        if (ext == EXT_MIRROR_SINGLE) {
            if (gdim == 1) g = 0;
            else {
                const int p = 2 * gdim - 2;
                g %= p;
                if (g < 0) g += p;
                if (g >= gdim) g = p - g;
            }
        } else {
            fill = (ext == EXT_CONST) ? ext_value : 0.f;
            return -1;
        }
    }
    const int l = g - gmin + off;
    if (l < 0 || l >= dim) { fill = 0.f; return -1; }  // outside the resident brick: never consumed
    return l;
}

template <int BX, int BY>
__global__ void __launch_bounds__(NT, 3) fwd_xy_kernel(Volume src, const Tile* __restrict__ tiles, float2* __restrict__ scratch,
                                                    size_t tile_stride, int bz, Twiddles tw) {
    using L = PlaneLayout<BX>;
    constexpr int H = BX / 2;
    constexpr int TX = H / 8;            // threads per row, 8 packed complex each
    constexpr int RPR = NT / TX;         // rows per round
    constexpr int RROUNDS = (BY + RPR - 1) / RPR;
    constexpr int TY = BY / 16;          // threads per column, 16 elements each
    constexpr int CPR = NT / TY;         // columns per round
    constexpr int CROUNDS = (H + CPR - 1) / CPR;
    extern __shared__ float2 plane[];

    const Tile tile = tiles[blockIdx.y];
    const int z = blockIdx.x;
    const int tid = threadIdx.x;
    const float2* __restrict__ twx = tw.t[ilog2<H>()];
    const float2* __restrict__ twbx = tw.t[ilog2<BX>()];
    const float2* __restrict__ twy = tw.t[ilog2<BY>()];

    float fillz = 0.f;
    const int zl = map_coord(tile.o[2] + z, src.dim[2], src.off[2], src.gmin[2], src.gdim[2], src.ext, src.ext_value, fillz);
    const int xo = tile.o[0] + src.off[0];  // array index of the window start
    const bool xfast = xo >= 0 && xo + BX <= src.dim[0];

    // ---- phase 1: rows, real -> half spectrum (packed: slot 0 = (X[0], X[H]))
#pragma unroll 1
    for (int rr = 0; rr < RROUNDS; ++rr) {
        const int row = rr * RPR + tid / TX;
        const int t = tid % TX;
        const bool valid = row < BY;
        float2 v[8];
        float fill = fillz;
        int yl = -1;
        if (valid && zl >= 0) yl = map_coord(tile.o[1] + row, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext, src.ext_value, fill);
        if (yl < 0) {
#pragma unroll
            for (int m = 0; m < 8; ++m) v[m] = make_float2(fill, fill);
        } else {
            const float* __restrict__ rowp = src.data + ((size_t)zl * src.dim[1] + yl) * (size_t)src.dim[0];
            if (xfast && ((reinterpret_cast<uintptr_t>(rowp + xo) & 7) == 0)) {
                const float2* __restrict__ p2 = reinterpret_cast<const float2*>(rowp + xo);
#pragma unroll
                for (int m = 0; m < 8; ++m) v[m] = __ldg(p2 + t + m * TX);
            } else {
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    float e[2];
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        float fx = 0.f;
                        const int xl = map_coord(tile.o[0] + 2 * (t + m * TX) + q, src.dim[0], src.off[0], src.gmin[0], src.gdim[0], src.ext,
                                                 src.ext_value, fx);
                        e[q] = xl >= 0 ? __ldg(rowp + xl) : fx;
                    }
                    v[m] = make_float2(e[0], e[1]);
                }
            }
        }
        auto put = [&](int idx, float2 val) { if (valid) plane[L::at(row, idx)] = val; };
        auto get = [&](int idx) { return valid ? plane[L::at(row, idx)] : make_float2(0.f, 0.f); };
        auto sync = [] { __syncwarp(); };
        line_fft<H, 8, false>(v, t, twx, put, get, sync);
        // untangle: X[k] (x2) = (Z[k] + conj Z[H-k]) - i w^k (Z[k] - conj Z[H-k])
        __syncwarp();
#pragma unroll
        for (int m = 0; m < 8; ++m) put(t + m * TX, v[m]);
        __syncwarp();
        float2 zp[8];
#pragma unroll
        for (int m = 0; m < 8; ++m) zp[m] = cconj(get((H - (t + m * TX)) & (H - 1)));
        __syncwarp();
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const int k = t + m * TX;
            float2 xk;
            if (k == 0) {
                xk = make_float2(2.f * (v[m].x + v[m].y), 2.f * (v[m].x - v[m].y));
            } else {
                const float2 s = cadd(v[m], zp[m]), d = csub(v[m], zp[m]);
                const float2 w = __ldg(&twbx[k]);
                const float2 wd = cmul(w, d);
                xk = make_float2(s.x + wd.y, s.y - wd.x);  // s - i*wd
            }
            put(k, xk);
        }
    }
    __syncthreads();

    // ---- phase 2: columns (H packed columns), write out H+1 unpacked columns
    float2* __restrict__ out = scratch + (size_t)blockIdx.y * tile_stride + (size_t)z * ((H + 1) * BY);
#pragma unroll 1
    for (int cr = 0; cr < CROUNDS; ++cr) {
        const int col = cr * CPR + tid % CPR;
        const int t = tid / CPR;
        const bool valid = col < H;
        float2 v[16];
#pragma unroll
        for (int m = 0; m < 16; ++m) v[m] = valid ? plane[L::at(t + m * TY, col)] : make_float2(0.f, 0.f);
        auto put = [&](int idx, float2 val) { if (valid) plane[L::at(idx, col)] = val; };
        auto get = [&](int idx) { return valid ? plane[L::at(idx, col)] : make_float2(0.f, 0.f); };
        auto sync = [] { __syncthreads(); };
        if constexpr (TY > 1) __syncthreads();  // everyone has gathered before the in-place scatter
        line_fft<BY, 16, false>(v, t, twy, put, get, sync);
        if (valid && col > 0) {
#pragma unroll
            for (int m = 0; m < 16; ++m) out[(t + m * TY) * (H + 1) + col] = v[m];
        }
        if (cr == 0) {
            // column 0 carries DC + i*Nyquist of x: split with the Hermitian partner in y
            __syncthreads();
            if (col == 0) {
#pragma unroll
                for (int m = 0; m < 16; ++m) plane[L::at(t + m * TY, 0)] = v[m];
            }
            __syncthreads();
            for (int ky = tid; ky < BY; ky += NT) {
                const float2 c = plane[L::at(ky, 0)];
                const float2 cm = cconj(plane[L::at((BY - ky) & (BY - 1), 0)]);
                const float2 s = cadd(c, cm), d = csub(c, cm);
                out[ky * (H + 1)] = make_float2(0.5f * s.x, 0.5f * s.y);
                out[ky * (H + 1) + H] = make_float2(0.5f * d.y, -0.5f * d.x);  // d / (2i)
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// z pass.  MODE 0: forward FFT only, scaled (builds a kernel spectrum).
//          MODE 1: forward FFT, multiply with the kernel spectrum, inverse FFT;
//                  only the planes later consumed by inv_xy are written back.
template <int BZ> struct ZCfg {
    static constexpr int EPT = BZ >= 16 ? 16 : BZ;
    static constexpr int TZ = BZ / EPT;
    static constexpr int PW = (256 / TZ) < 16 ? 16 : (256 / TZ);
    static constexpr int THREADS = PW * TZ;
    static constexpr size_t SMEM = (size_t)PW * BZ * sizeof(float2);
    static constexpr int MIN_CTAS = THREADS >= 512 ? 2 : 3;
};

template <int BZ, int MODE>
__global__ void __launch_bounds__(ZCfg<BZ>::THREADS, ZCfg<BZ>::MIN_CTAS) z_pass_kernel(float2* __restrict__ scratch, size_t tile_stride,
                                                                   const float2* __restrict__ kspec, const Tile* __restrict__ tiles,
                                                                   int plane_elems, float scale, Twiddles tw) {
    using C = ZCfg<BZ>;
    constexpr int EPT = C::EPT, TZ = C::TZ, PW = C::PW;
    extern __shared__ float2 xch[];
    const int tid = threadIdx.x;
    const int p = tid % PW, t = tid / PW;
    const int pos = blockIdx.x * PW + p;
    const bool valid = pos < plane_elems;
    const float2* __restrict__ twz = tw.t[ilog2<BZ>()];
    float2* __restrict__ base = scratch + (size_t)blockIdx.y * tile_stride + pos;

    float2 v[EPT];
#pragma unroll
    for (int m = 0; m < EPT; ++m) v[m] = valid ? base[(size_t)(t + m * TZ) * plane_elems] : make_float2(0.f, 0.f);

    auto put = [&](int idx, float2 val) { xch[idx * PW + p] = val; };
    auto get = [&](int idx) { return xch[idx * PW + p]; };
    auto sync = [] { __syncthreads(); };
    line_fft<BZ, EPT, false>(v, t, twz, put, get, sync);

    if constexpr (MODE == 0) {
        if (valid) {
#pragma unroll
            for (int m = 0; m < EPT; ++m) base[(size_t)(t + m * TZ) * plane_elems] = make_float2(v[m].x * scale, v[m].y * scale);
        }
    } else {
        const float2* __restrict__ kb = kspec + pos;
#pragma unroll
        for (int m = 0; m < EPT; ++m) {
            const float2 k = valid ? __ldg(kb + (size_t)(t + m * TZ) * plane_elems) : make_float2(0.f, 0.f);
            v[m] = cmul(v[m], k);
        }
        if constexpr (TZ > 1) __syncthreads();
        line_fft<BZ, EPT, true>(v, t, twz, put, get, sync);
        const Tile tile = tiles[blockIdx.y];
        const int zlo = tile.e[2] - tile.o[2], zhi = zlo + tile.n[2];
        if (valid) {
#pragma unroll
            for (int m = 0; m < EPT; ++m) {
                const int z = t + m * TZ;
                if (z >= zlo && z < zhi) base[(size_t)z * plane_elems] = v[m];
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// per-voxel epilogues
__device__ __forceinline__ float quotient_value(float img, float blurred) {
    // MVDeconvolution.java:496-499 : imgValue > 0 ? imgValue / psiBlurred : 1
    return img > 0.f ? __fdiv_rn(img, blurred) : 1.f;
}

__device__ __forceinline__ float next_psi_value(float last, float integral, float weight, double lambda, float min_value) {
    // MVDeconvolution.java:671-703 (computeNextValue) and :705 (tikhonov); no FMA contraction
    const float value = __fmul_rn(last, integral);
    float adjusted;
    if (value > 0.f) {
        if (lambda > 0.0) {
            const double s = sqrt(__dadd_rn(1.0, __dmul_rn(__dmul_rn(2.0, lambda), (double)value)));
            adjusted = (float)(__ddiv_rn(__dsub_rn(s, 1.0), lambda));
        } else {
            adjusted = value;
        }
    } else {
        adjusted = min_value;
    }
    const float next = isnan(adjusted) ? min_value : fmaxf(min_value, adjusted);
    return __fadd_rn(last, __fmul_rn(__fsub_rn(next, last), weight));
}

template <int BX, int BY, int EPI>
__global__ void __launch_bounds__(NT, 3) inv_xy_kernel(const float2* __restrict__ scratch, size_t tile_stride, const Tile* __restrict__ tiles,
                                                    EpilogueArgs ep, Twiddles tw) {
    using L = PlaneLayout<BX>;
    constexpr int H = BX / 2;
    constexpr int TX = H / 8;
    constexpr int RPR = NT / TX;
    constexpr int TY = BY / 16;
    constexpr int CPR = NT / TY;
    constexpr int CROUNDS = (H + CPR - 1) / CPR;
    extern __shared__ float2 plane[];
    __shared__ double s_sum[NT / 32];
    __shared__ float s_max[NT / 32];

    const Tile tile = tiles[blockIdx.y];
    if ((int)blockIdx.x >= tile.n[2]) return;
    const int zt = tile.e[2] - tile.o[2] + blockIdx.x;  // plane inside the tile
    const int zv = tile.e[2] + blockIdx.x;              // plane inside the volume
    const int tid = threadIdx.x;
    const float2* __restrict__ twx = tw.t[ilog2<H>()];
    const float2* __restrict__ twbx = tw.t[ilog2<BX>()];
    const float2* __restrict__ twy = tw.t[ilog2<BY>()];
    const float2* __restrict__ in = scratch + (size_t)blockIdx.y * tile_stride + (size_t)zt * ((H + 1) * BY);

    // ---- phase 1: inverse columns; column 0 re-packs DC + i*Nyquist
#pragma unroll 1
    for (int cr = 0; cr < CROUNDS; ++cr) {
        const int col = cr * CPR + tid % CPR;
        const int t = tid / CPR;
        const bool valid = col < H;
        float2 v[16];
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            const int ky = t + m * TY;
            float2 a = valid ? in[ky * (H + 1) + col] : make_float2(0.f, 0.f);
            if (col == 0) {
                const float2 b = in[ky * (H + 1) + H];
                a = make_float2(a.x - b.y, a.y + b.x);
            }
            v[m] = a;
        }
        auto put = [&](int idx, float2 val) { if (valid) plane[L::at(idx, col)] = val; };
        auto get = [&](int idx) { return valid ? plane[L::at(idx, col)] : make_float2(0.f, 0.f); };
        auto sync = [] { __syncthreads(); };
        line_fft<BY, 16, true>(v, t, twy, put, get, sync);
        if constexpr (TY > 1) __syncthreads();
#pragma unroll
        for (int m = 0; m < 16; ++m) put(t + m * TY, v[m]);
    }
    __syncthreads();

    // ---- phase 2: inverse rows (C2R through the packed half-length transform) + epilogue.
    // Only the rows of the effective region are transformed; the others would be discarded.
    double lsum = 0.0;
    float lmax = 0.f;
    const int lx = tile.e[0] - tile.o[0], ly = tile.e[1] - tile.o[1];
    const int n0 = tile.n[0], n1 = tile.n[1];
    const int rrounds = (n1 + RPR - 1) / RPR;
#pragma unroll 1
    for (int rr = 0; rr < rrounds; ++rr) {
        const int yl = rr * RPR + tid / TX;  // row inside the effective region
        const int row = ly + yl;
        const int t = tid % TX;
        const bool valid = yl < n1;
        float2 v[8];
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const int k = t + m * TX;
            float2 zk = make_float2(0.f, 0.f);
            if (valid) {
                const float2 xk = plane[L::at(row, k)];
                if (k == 0) {
                    zk = make_float2(xk.x + xk.y, xk.x - xk.y);
                } else {
                    const float2 xp = cconj(plane[L::at(row, H - k)]);
                    const float2 s = cadd(xk, xp), d = csub(xk, xp);
                    const float2 w = cconj(__ldg(&twbx[k]));
                    const float2 wd = cmul(w, d);
                    zk = make_float2(s.x - wd.y, s.y + wd.x);  // s + i*wd
                }
            }
            v[m] = zk;
        }
        __syncwarp();
        auto put = [&](int idx, float2 val) { if (valid) plane[L::at(row, idx)] = val; };
        auto get = [&](int idx) { return valid ? plane[L::at(row, idx)] : make_float2(0.f, 0.f); };
        auto sync = [] { __syncwarp(); };
        line_fft<H, 8, true>(v, t, twx, put, get, sync);

        if (valid) {
            const size_t rowbase = ((size_t)(zv + ep.out_off[2]) * ep.out_dim[1] + (tile.e[1] + yl + ep.out_off[1])) * (size_t)ep.out_dim[0] +
                                   tile.e[0] + ep.out_off[0];
            const size_t auxbase = ((size_t)zv * ep.aux_dim[1] + (tile.e[1] + yl)) * (size_t)ep.aux_dim[0] + tile.e[0];
            // thread owns the pairs x = 2k-lx, 2k-lx+1; lx is a multiple of 4, so a pair is 8-byte
            // aligned whenever the row start is
            const bool vec = ((rowbase | auxbase) & 1) == 0;
            float2 a[8], b[8];
            unsigned full = 0, part = 0;
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                const int xl = 2 * (t + m * TX) - lx;
                if (xl >= 0 && xl + 1 < n0) full |= 1u << m;
                else if (xl + 1 == 0 || xl == n0 - 1) part |= 1u << m;
            }
            if (vec) {
                // issue every load before the first use
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    const int xl = 2 * (t + m * TX) - lx;
                    a[m] = make_float2(0.f, 0.f);
                    b[m] = make_float2(ep.weight_const, ep.weight_const);
                    if (full >> m & 1) {
                        if constexpr (EPI == EPI_QUOTIENT) a[m] = __ldg(reinterpret_cast<const float2*>(ep.img + auxbase + xl));
                        if constexpr (EPI == EPI_UPDATE) {
                            a[m] = *reinterpret_cast<const float2*>(ep.out + rowbase + xl);
                            if (ep.weight) b[m] = __ldg(reinterpret_cast<const float2*>(ep.weight + auxbase + xl));
                        }
                    }
                }
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    if (!(full >> m & 1)) continue;
                    const int xl = 2 * (t + m * TX) - lx;
                    const float c0 = v[m].x + ep.add_const, c1 = v[m].y + ep.add_const;
                    float2 r;
                    if constexpr (EPI == EPI_STORE) r = make_float2(c0, c1);
                    else if constexpr (EPI == EPI_QUOTIENT) r = make_float2(quotient_value(a[m].x, c0), quotient_value(a[m].y, c1));
                    else {
                        r.x = next_psi_value(a[m].x, c0, b[m].x, ep.lambda, ep.min_value);
                        r.y = next_psi_value(a[m].y, c1, b[m].y, ep.lambda, ep.min_value);
                        const float ch0 = fabsf(__fsub_rn(r.x, a[m].x)), ch1 = fabsf(__fsub_rn(r.y, a[m].y));
                        lsum += (double)ch0 + (double)ch1;
                        lmax = fmaxf(lmax, fmaxf(ch0, ch1));  // fmaxf drops NaN
                    }
                    *reinterpret_cast<float2*>(ep.out + rowbase + xl) = r;
                }
            } else {
                part |= full;
            }
            // scalar path: pairs straddling the region border, or unaligned rows
            if (part) {
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    if (!(part >> m & 1)) continue;
                    const int xl = 2 * (t + m * TX) - lx;
                    const float c[2] = {v[m].x + ep.add_const, v[m].y + ep.add_const};
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int x = xl + q;
                        if (x < 0 || x >= n0) continue;
                        const size_t idx = rowbase + x, aidx = auxbase + x;
                        if constexpr (EPI == EPI_STORE) {
                            ep.out[idx] = c[q];
                        } else if constexpr (EPI == EPI_QUOTIENT) {
                            ep.out[idx] = quotient_value(__ldg(ep.img + aidx), c[q]);
                        } else {
                            const float last = ep.out[idx];
                            const float w = ep.weight ? __ldg(ep.weight + aidx) : ep.weight_const;
                            const float nv = next_psi_value(last, c[q], w, ep.lambda, ep.min_value);
                            ep.out[idx] = nv;
                            const float ch = fabsf(__fsub_rn(nv, last));
                            lsum += (double)ch;
                            lmax = fmaxf(lmax, ch);
                        }
                    }
                }
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
                double s = 0.0; float mx = 0.f;
                for (int i = 0; i < NT / 32; ++i) { s += s_sum[i]; mx = fmaxf(mx, s_max[i]); }
                atomicAdd(ep.sum_change, s);
                atomicMax(ep.max_change_bits, __float_as_uint(mx));
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// host launchers
bool xy_size_supported(int b) { return b == 32 || b == 64 || b == 128; }
bool z_size_supported(int b) { return b == 16 || b == 32 || b == 64 || b == 128 || b == 256 || b == 512; }
size_t plane_elems(int bx, int by) { return (size_t)(bx / 2 + 1) * by; }
size_t tile_scratch_elems(int bx, int by, int bz) { return plane_elems(bx, by) * bz; }

static inline Twiddles to_dev(const TwiddleTables& t) {
    Twiddles d;
    for (int i = 0; i < 11; ++i) d.t[i] = t.t[i];
    return d;
}
template <int BX> static constexpr size_t xy_smem(int by) { return (size_t)by * (BX / 2 + 8) * sizeof(float2); }

#define MVD_XY_DISPATCH(bx, by, F)                                                            \
    do {                                                                                      \
        if (bx == 32 && by == 32) { F(32, 32); } else if (bx == 32 && by == 64) { F(32, 64); } \
        else if (bx == 32 && by == 128) { F(32, 128); } else if (bx == 64 && by == 32) { F(64, 32); } \
        else if (bx == 64 && by == 64) { F(64, 64); } else if (bx == 64 && by == 128) { F(64, 128); } \
        else if (bx == 128 && by == 32) { F(128, 32); } else if (bx == 128 && by == 64) { F(128, 64); } \
        else if (bx == 128 && by == 128) { F(128, 128); } else return cudaErrorInvalidValue;      \
    } while (0)

#define MVD_Z_DISPATCH(bz, F)                                                               \
    do {                                                                                    \
        switch (bz) {                                                                       \
            case 16: F(16); break; case 32: F(32); break; case 64: F(64); break;           \
            case 128: F(128); break; case 256: F(256); break; case 512: F(512); break;     \
            default: return cudaErrorInvalidValue;                                          \
        }                                                                                   \
    } while (0)

cudaError_t configure_conv_kernels() {
    cudaError_t e = cudaSuccess;
#define CFG_XY(BX, BY)                                                                                                           \
    {                                                                                                                            \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(fwd_xy_kernel<BX, BY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xy_smem<BX>(BY)); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xy_smem<BX>(BY)); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_QUOTIENT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xy_smem<BX>(BY)); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_UPDATE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xy_smem<BX>(BY)); \
    }
    CFG_XY(32, 32) CFG_XY(32, 64) CFG_XY(32, 128) CFG_XY(64, 32) CFG_XY(64, 64) CFG_XY(64, 128) CFG_XY(128, 32) CFG_XY(128, 64) CFG_XY(128, 128)
#undef CFG_XY
#define CFG_Z(BZ)                                                                                                                 \
    {                                                                                                                             \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(z_pass_kernel<BZ, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZCfg<BZ>::SMEM); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(z_pass_kernel<BZ, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZCfg<BZ>::SMEM); \
    }
    CFG_Z(16) CFG_Z(32) CFG_Z(64) CFG_Z(128) CFG_Z(256) CFG_Z(512)
#undef CFG_Z
    return e;
}

cudaError_t launch_fwd_xy(int bx, int by, int bz, const Volume& src, const Tile* d_tiles, int n_tiles, float2* scratch,
                          const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0) return cudaSuccess;
    const size_t stride = tile_scratch_elems(bx, by, bz);
    const dim3 grid(bz, n_tiles);
#define F(BX, BY) fwd_xy_kernel<BX, BY><<<grid, NT, xy_smem<BX>(BY), st>>>(src, d_tiles, scratch, stride, bz, to_dev(tw))
    MVD_XY_DISPATCH(bx, by, F);
#undef F
    return cudaGetLastError();
}

cudaError_t launch_z_pass(int bx, int by, int bz, int mode, float2* scratch, const float2* kspec, const Tile* d_tiles, int n_tiles,
                          float scale, const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0) return cudaSuccess;
    const size_t stride = tile_scratch_elems(bx, by, bz);
    const int pe = (int)plane_elems(bx, by);
#define F(BZ)                                                                                                          \
    {                                                                                                                  \
        const dim3 grid((pe + ZCfg<BZ>::PW - 1) / ZCfg<BZ>::PW, n_tiles);                                              \
        if (mode == 0) z_pass_kernel<BZ, 0><<<grid, ZCfg<BZ>::THREADS, ZCfg<BZ>::SMEM, st>>>(scratch, stride, kspec, d_tiles, pe, scale, to_dev(tw)); \
        else z_pass_kernel<BZ, 1><<<grid, ZCfg<BZ>::THREADS, ZCfg<BZ>::SMEM, st>>>(scratch, stride, kspec, d_tiles, pe, scale, to_dev(tw)); \
    }
    MVD_Z_DISPATCH(bz, F);
#undef F
    return cudaGetLastError();
}

cudaError_t launch_inv_xy(int bx, int by, int bz, int epilogue, const float2* scratch, const Tile* d_tiles, int n_tiles, int max_nz,
                          const EpilogueArgs& ep, const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0 || max_nz <= 0) return cudaSuccess;
    const size_t stride = tile_scratch_elems(bx, by, bz);
    const dim3 grid(max_nz, n_tiles);
#define F(BX, BY)                                                                                                        \
    {                                                                                                                    \
        if (epilogue == EPI_STORE) inv_xy_kernel<BX, BY, EPI_STORE><<<grid, NT, xy_smem<BX>(BY), st>>>(scratch, stride, d_tiles, ep, to_dev(tw)); \
        else if (epilogue == EPI_QUOTIENT) inv_xy_kernel<BX, BY, EPI_QUOTIENT><<<grid, NT, xy_smem<BX>(BY), st>>>(scratch, stride, d_tiles, ep, to_dev(tw)); \
        else inv_xy_kernel<BX, BY, EPI_UPDATE><<<grid, NT, xy_smem<BX>(BY), st>>>(scratch, stride, d_tiles, ep, to_dev(tw)); \
    }
    MVD_XY_DISPATCH(bx, by, F);
#undef F
    return cudaGetLastError();
}

}  // namespace mvd
