// Tiled overlap-save 3-D FFT convolution for sm_100a: three kernels per tile batch.
//
//   fwd_xy : source volume (+ on-the-fly out-of-bounds rule) -> 2-D R2C FFT of z-planes of a
//            tile in shared memory -> half-spectrum planes in scratch
//   z_pass : forward z FFT, pointwise multiply with the resident kernel spectrum,
//            inverse z FFT, in place on PW plane positions x BZ per CTA
//   inv_xy : inverse 2-D C2R FFT of the effective z-planes in shared memory with the
//            fused per-voxel epilogue (store / quotient / multiplicative update)
//
// Replaces, for one block, copyBlock -> FFTConvolution/convolution3DfftCUDAInPlace ->
// pasteBlock of the reference (MVDeconFFTThreads.java:18-94, Block.java:114-359,
// J/fftMethods/FFTConvolution.java:470-551) and the elementwise passes
// MVDeconvolution.java:473-525 and :582-703.
//
// The kernels are instruction-issue bound before they are HBM bound (profiles/), so the code is
// organised to keep everything that is not a butterfly out of the per-plane loop: out-of-bounds
// index maps, twiddle tables and tile geometry are set up once per CTA in shared memory, the
// shared-memory plane uses a padded linear layout (one pad slot per 8) that is bank-conflict free
// for the strided Stockham scatters without per-access swizzle arithmetic, column twiddles come
// from the constant bank (warp-uniform), and all register arrays are statically indexed.
#include "conv_kernels.h"
#include <climits>
#include "fft_device.cuh"
#include "row128.cuh"
#include "z_items.cuh"
#include "device_common.cuh"
#include "inv_rows128.cuh"

#include <algorithm>

namespace mvd {

// threads per CTA of the xy kernels: 256, two or three CTAs per SM; windows with a 256-point side hold a plane of
// > 113 KB, i.e. one CTA per SM, and run 512 threads
template <int BX, int BY> struct XYThreads { static constexpr int NT = (BX > 128 || BY > 128) ? 512 : 256; };
#ifndef MVD_XY_ZPC
#define MVD_XY_ZPC 4  // planes per work item of the xy kernels (measured: 1 -> 32.1, 2 -> 31.9, 4 -> 31.5 ms per iteration)
#endif
#ifndef MVD_PREFETCH
#define MVD_PREFETCH 0
#endif
static constexpr int XY_ZPC = MVD_XY_ZPC;  // z-planes per CTA of the xy kernels

// exp(-2 pi i n / 256): column twiddles of the xy kernels (the index is warp-uniform there)
__constant__ float2 c_tw256[256];

struct ConstTw256 {
    int shift;  // log2(256 / N)
    __device__ __forceinline__ float2 operator[](int i) const { return c_tw256[i << shift]; }
};

// shared-memory plane: BY rows of H = BX/2 packed complex columns.
// Generic sizes: element k of a row lives in slot k + (k >> 3) and the row pitch is P = H + H/8,
// i.e. 16 (H=64), 8 (H=32), 4 (H=16) banks modulo 32 -- rows of a warp never collide and a
// stride-8 scatter becomes a stride-9 one.
// BX = 128 (the hot size): natural slots, pitch 68 (= 4 slots modulo 16): the fast row passes put
// 4 threads on a row and 8 rows in a warp, so the 4 rows of a half-warp sit 4 slots apart and
// every access pattern below covers the 16 slots (32 banks) exactly once; the column passes read
// 32 consecutive columns per warp, conflict-free in natural order.
template <int BX> struct PlaneLayout {
    static constexpr int H = BX / 2;
    static constexpr int P = BX == 128 ? 68 : H + H / 8;
    __device__ __forceinline__ static constexpr int slot(int k) { return BX == 128 ? k : k + (k >> 3); }
};

// identity the optimiser cannot see through (keeps it from sharing address arithmetic across phases)
__device__ __forceinline__ int opaque(int x) {
    asm volatile("mov.b32 %0, %0;" : "+r"(x));
    return x;
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <int BX, int BY, int NT_> struct XYCfgT {
    using L = PlaneLayout<BX>;
    static constexpr int NT = NT_;
    static constexpr int H = BX / 2;
    static constexpr int TX = H / 8;            // threads per row, 8 packed complex each
    static constexpr int RPR = NT / TX;         // rows per round
    static constexpr int TY = BY / 16;          // threads per column, 16 elements each
    static constexpr int CPR = NT / TY;         // columns per round
    static constexpr int CROUNDS = (H + CPR - 1) / CPR;
    static constexpr bool CFULL = (H % CPR) == 0;
    static constexpr size_t PLANE = (size_t)BY * L::P * sizeof(float2);
    // forward: plane | row twiddles (H) | untangle twiddles (H) | column 0 staging (BY) | x map (BX)
    // (BX = 128: the two twiddle areas hold the per-thread tables of the fast row pass, 64 + 32 float2)
    static constexpr size_t SMEM_FWD = PLANE + (size_t)(2 * H + BY) * sizeof(float2) + (size_t)BX * sizeof(int);
    // inverse: plane | row twiddles (H) | untangle twiddles (H)
    static constexpr size_t SMEM_INV = PLANE + (size_t)(2 * H) * sizeof(float2);
    static_assert(BX != 128 || 2 * H >= 96, "fast-row tables need 96 float2");
    // 228 KB per SM, 1 KB reserved per resident CTA (+ the inverse kernel's static reduction arrays)
    static constexpr int MIN_CTAS = NT > 256 ? 1 : ((SMEM_FWD + 1024 + 128) * 3 <= 233472 ? 3 : ((SMEM_FWD + 1024 + 128) * 2 <= 233472 ? 2 : 1));
};
#ifndef MVD_INV_MINCTAS
#define MVD_INV_MINCTAS 2  // 2 x 256 threads at <= 128 registers beat 3 x 256 at 80 with spills (measured)
#endif
template <int BX, int BY> using XYCfg = XYCfgT<BX, BY, XYThreads<BX, BY>::NT>;
// the inverse kernel's own thread count: MVD_INV_NT128 = three 128-thread CTAs per SM for 128 x 128 windows (experiment)
#ifndef MVD_INV_NT128
#define MVD_INV_NT128 1  // measured: inv_xy 11.17 -> 10.99 ms per four convolutions (168 registers, 20 B of spills instead of 176)
#endif
template <int BX, int BY> struct InvThreads {
    static constexpr int NT = (MVD_INV_NT128 != 0 && BX == 128 && BY == 128) ? 128 : XYThreads<BX, BY>::NT;
    static constexpr int MIN_CTAS_WANT = NT == 128 ? 3 : MVD_INV_MINCTAS;
};
template <int BX, int BY> using InvCfg = XYCfgT<BX, BY, InvThreads<BX, BY>::NT>;

// ---------------------------------------------------------------------------------------------
#ifndef MVD_FWD_MINCTAS
#define MVD_FWD_MINCTAS 3
#endif
#ifndef MVD_FWD_PIPE
#define MVD_FWD_PIPE 1  // 128 x 128 windows: both row rounds of a plane requested up front (8.59 vs 9.09 ms per 4 convolutions)
#endif
// the up-front loads of the 128 x 128 kernel need 126 registers: two CTAs per SM (measured equal to three at 77 registers without them)
template <int BX, int BY> struct FwdCfg {
    static constexpr bool PIPE = MVD_FWD_PIPE != 0 && BX == 128 && BY == 2 * (XYThreads<BX, BY>::NT / 4);
    static constexpr int WANT = PIPE ? 2 : MVD_FWD_MINCTAS;
    static constexpr int MIN_CTAS = XYCfg<BX, BY>::MIN_CTAS < WANT ? XYCfg<BX, BY>::MIN_CTAS : WANT;
};
template <int BX, int BY>
__global__ void __launch_bounds__((XYThreads<BX, BY>::NT), FwdCfg<BX, BY>::MIN_CTAS) fwd_xy_kernel(Volume src, const Tile* __restrict__ tiles,
                                                                           float2* __restrict__ scratch, size_t tile_stride, int bz,
                                                                           int n_items, Twiddles tw) {
    using C = XYCfg<BX, BY>;
    using L = PlaneLayout<BX>;
    constexpr int NT = C::NT;
    constexpr int H = C::H, P = L::P, TX = C::TX, RPR = C::RPR, TY = C::TY, CPR = C::CPR;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* plane = reinterpret_cast<float2*>(smem_raw);
    float2* twr = plane + (size_t)BY * P;   // exp(-2 pi i n / H)
    float2* twu = twr + H;                  // exp(-2 pi i n / BX), n < H
    float2* col0 = twu + H;                 // staging of column 0 (DC + i Nyquist)
    int* xmap = reinterpret_cast<int*>(col0 + BY);

    const int tid = threadIdx.x;
    // ---- per-CTA setup
    if constexpr (BX == 128) {
        row128_fill_tables(twr, twu, tw.t[6], tw.t[7], false, tid, NT);
    } else {
        for (int i = tid; i < H; i += NT) {
            twr[i] = __ldg(&tw.t[ilog2<H>()][i]);
            twu[i] = __ldg(&tw.t[ilog2<BX>()][i]);
        }
    }
    const float fillv = src.ext == EXT_CONST ? src.ext_value : 0.f;
    const int chunks = (bz + XY_ZPC - 1) / XY_ZPC;
    // work items = (tile, chunk of XY_ZPC planes), chunk fastest; the grid may be smaller than the
    // item count (persistent CTAs) so that CTAs of other kernels fit beside this one on the SM
#pragma unroll 1
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int ti = item / chunks;
    const Tile tile = tiles[ti];
    const int z0 = (item - ti * chunks) * XY_ZPC;
    __syncthreads();  // the previous item is done with xmap and the plane
    for (int i = tid; i < BX; i += NT) xmap[i] = map_coord(tile.o[0] + i, src.dim[0], src.off[0], src.gmin[0], src.gdim[0], src.ext);
    const int xo = tile.o[0] + src.off[0];  // array index of the window start
    const int xg = tile.o[0] + src.gmin[0];
    // whole window inside the array and inside the global domain, rows 8-byte aligned
    const bool xfast = xo >= 0 && xo + BX <= src.dim[0] && xg >= 0 && xg + BX <= src.gdim[0] && ((xo | src.dim[0]) & 1) == 0 &&
                       (reinterpret_cast<uintptr_t>(src.data) & 7) == 0;
    __syncthreads();

    // row phase thread mapping
    const int rt = tid % TX;
    const int rrow = tid / TX;
    // column phase thread mapping
    const int cc = tid % CPR;
    const int cu = tid / CPR;
    const ConstTw256 twc{8 - ilog2<BY>()};

    // the same for every plane of the item (the compiler recomputed them per plane: 134 integer instructions, 6 % of the
    // kernel): which quads of a row are plain float4 loads, and where the thread's two rows of the window lie in the array
    unsigned qmask = 0;
    int ymap0 = 0, ymap1 = 0;
    if constexpr (BX == 128) {
        const int q = tid & 3;
        const bool al4 = ((xo | src.dim[0]) & 3) == 0 && (reinterpret_cast<uintptr_t>(src.data) & 15) == 0;
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const int xa = xo + 16 * m + 4 * q, xgl = xg + 16 * m + 4 * q;
            if (al4 && xa >= 0 && xa + 3 < src.dim[0] && xgl >= 0 && xgl + 3 < src.gdim[0]) qmask |= 1u << m;
        }
        if constexpr (FwdCfg<BX, BY>::PIPE) {
            ymap0 = map_coord(tile.o[1] + (tid >> 2), src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
            ymap1 = map_coord(tile.o[1] + (tid >> 2) + NT / 4, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
        }
    }

#pragma unroll 1
    for (int zi = 0; zi < XY_ZPC; ++zi) {
        const int z = z0 + zi;
        if (z >= bz) break;
        const int zl = map_coord(tile.o[2] + z, src.dim[2], src.off[2], src.gmin[2], src.gdim[2], src.ext);
        const float* __restrict__ zbase = src.data + (size_t)(zl < 0 ? 0 : zl) * src.dim[1] * (size_t)src.dim[0];

        // ---- phase 1: rows, real -> half spectrum (packed: slot 0 = (X[0], X[H]))
        if constexpr (BX == 128) {
            const int q = tid & 3;
            if ((MVD_PREFETCH & 1) && xfast && zi + 1 < XY_ZPC && z + 1 < bz) {
                // pull the next plane's rows towards L2 while this plane is transformed
                const int zn = map_coord(tile.o[2] + z + 1, src.dim[2], src.off[2], src.gmin[2], src.gdim[2], src.ext);
                if (zn >= 0) {
                    for (int row = tid >> 2; row < BY; row += NT / 4) {
                        const int yn = map_coord(tile.o[1] + row, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
                        if (yn >= 0) prefetch_l2(src.data + ((size_t)zn * src.dim[1] + yn) * (size_t)src.dim[0] + xo + 32 * q);
                    }
                }
            }
            // quads (4 consecutive samples) of this thread that lie inside the array and the global
            // domain are loaded as float4 even in tiles that touch the volume border
            auto load_row = [&](int row, int ymapped, float2 (&a)[8], float2 (&b)[8]) {
                const int yl = zl < 0 ? zl : (ymapped != INT_MIN ? ymapped : map_coord(tile.o[1] + row, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext));
                if (yl < 0) {
                    const float f = yl == -1 ? fillv : 0.f;
#pragma unroll
                    for (int m = 0; m < 8; ++m) a[m] = b[m] = make_float2(f, f);
                } else {
                    const float* __restrict__ rowp = zbase + (size_t)yl * src.dim[0];
                    const float4* __restrict__ p4 = reinterpret_cast<const float4*>(rowp + xo) + q;
#pragma unroll
                    for (int m = 0; m < 8; ++m) {
                        if (qmask >> m & 1) {
                            const float4 f = __ldg(p4 + 4 * m);
                            a[m] = make_float2(f.x, f.y);
                            b[m] = make_float2(f.z, f.w);
                        } else {
                            float e[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const int xi = xmap[16 * m + 4 * q + i];
                                e[i] = xi >= 0 ? __ldg(rowp + xi) : (xi == -1 ? fillv : 0.f);
                            }
                            a[m] = make_float2(e[0], e[1]);
                            b[m] = make_float2(e[2], e[3]);
                        }
                    }
                }
            };
            if constexpr (FwdCfg<BX, BY>::PIPE) {
                // both row rounds of the plane requested up front: one exposed load latency per plane instead of two
                float2 a0[8], b0[8], a1[8], b1[8];
                const int r0 = tid >> 2, r1 = r0 + NT / 4;
                load_row(r0, ymap0, a0, b0);
                load_row(r1, ymap1, a1, b1);
                row128_forward(a0, b0, plane + r0 * P, twr, twu, q, 0xffffffffu);
                row128_forward(a1, b1, plane + r1 * P, twr, twu, q, 0xffffffffu);
            } else {
#pragma unroll 1
                for (int row = tid >> 2; row < BY; row += NT / 4) {
                    float2 a[8], b[8];
                    load_row(row, INT_MIN, a, b);
                    row128_forward(a, b, plane + row * P, twr, twu, q, 0xffffffffu);
                }
            }
        } else {
#pragma unroll 1
        for (int row = rrow; row < BY; row += RPR) {
            float2 v[8];
            const int yl = zl < 0 ? zl : map_coord(tile.o[1] + row, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
            if (yl < 0) {
                const float f = yl == -1 ? fillv : 0.f;
#pragma unroll
                for (int m = 0; m < 8; ++m) v[m] = make_float2(f, f);
            } else {
                const float* __restrict__ rowp = zbase + (size_t)yl * src.dim[0];
                if (xfast) {
                    const float2* __restrict__ p2 = reinterpret_cast<const float2*>(rowp + xo) + rt;
#pragma unroll
                    for (int m = 0; m < 8; ++m) v[m] = __ldg(p2 + m * TX);
                } else {
#pragma unroll
                    for (int m = 0; m < 8; ++m) {
                        const int x0 = xmap[2 * (rt + m * TX)], x1 = xmap[2 * (rt + m * TX) + 1];
                        v[m].x = x0 >= 0 ? __ldg(rowp + x0) : (x0 == -1 ? fillv : 0.f);
                        v[m].y = x1 >= 0 ? __ldg(rowp + x1) : (x1 == -1 ? fillv : 0.f);
                    }
                }
            }
            float2* __restrict__ rp = plane + row * P;
            auto put = [&](int idx, float2 val) { rp[L::slot(idx)] = val; };
            auto get = [&](int idx) { return rp[L::slot(idx)]; };
            auto sync = [] { __syncwarp(); };
            line_fft<H, 8, false>(v, rt, twr, put, get, sync);
            // untangle: X[k] (x2) = (Z[k] + conj Z[H-k]) - i w^k (Z[k] - conj Z[H-k])
            __syncwarp();
#pragma unroll
            for (int m = 0; m < 8; ++m) rp[rt + m * TX + ((rt + m * TX) >> 3)] = v[m];
            __syncwarp();
            float2 zp[8];
#pragma unroll
            for (int m = 0; m < 8; ++m) zp[m] = cconj(get((H - (rt + m * TX)) & (H - 1)));
            __syncwarp();
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                const int k = rt + m * TX;
                const float2 s = cadd(v[m], zp[m]), d = csub(v[m], zp[m]);
                const float2 wd = cmul(twu[k], d);
                float2 xk = make_float2(s.x + wd.y, s.y - wd.x);  // s - i*wd
                if (m == 0 && rt == 0) xk = make_float2(2.f * (v[0].x + v[0].y), 2.f * (v[0].x - v[0].y));
                rp[L::slot(k)] = xk;
            }
        }
        }
        __syncthreads();

        // ---- phase 2: columns (H packed columns), write out H unpacked columns + the Nyquist column
        float2* __restrict__ out = scratch + (size_t)ti * tile_stride + (size_t)z * ((H + 1) * BY);
#pragma unroll 1
        for (int cr = 0; cr < C::CROUNDS; ++cr) {
            const int col = cr * CPR + cc;
            const bool valid = C::CFULL || col < H;
            float2* __restrict__ cp = plane + L::slot(valid ? col : 0);
            float2 v[16];
#pragma unroll
            for (int m = 0; m < 16; ++m) v[m] = cp[(cu + m * TY) * P];
            auto put = [&](int idx, float2 val) { if (valid) cp[idx * P] = val; };
            auto get = [&](int idx) { return cp[idx * P]; };
            auto sync = [] { __syncthreads(); };
            if constexpr (TY > 1) __syncthreads();  // everyone has gathered before the in-place scatter
            line_fft<BY, 16, false>(v, cu, twc, put, get, sync);
            if (valid && col > 0) {
#pragma unroll
                for (int m = 0; m < 16; ++m) out[(cu + m * TY) * H + col] = v[m];
            }
            if (cr == 0 && col == 0) {
#pragma unroll
                for (int m = 0; m < 16; ++m) col0[cu + m * TY] = v[m];
            }
        }
        __syncthreads();
        // column 0 carries DC + i*Nyquist of x: split with the Hermitian partner in y
        for (int ky = tid; ky < BY; ky += NT) {
            const float2 c = col0[ky];
            const float2 cm = cconj(col0[(BY - ky) & (BY - 1)]);
            const float2 s = cadd(c, cm), d = csub(c, cm);
            out[ky * H] = make_float2(0.5f * s.x, 0.5f * s.y);
            out[H * BY + ky] = make_float2(0.5f * d.y, -0.5f * d.x);  // d / (2i)
        }
        // the next plane's row phase rewrites the plane: all column gathers are complete (barrier above)
    }
    }  // items
}

// ---------------------------------------------------------------------------------------------
// z pass.  MODE 0: forward FFT only, scaled (builds a kernel spectrum).
//          MODE 1: forward FFT, multiply with the kernel spectrum, inverse FFT;
//                  only the planes later consumed by inv_xy are written back.
// Lanes run across plane positions (coalesced 8-byte accesses, PW*8 contiguous bytes per
// z index); a z line is shared by TZ threads with EPT elements each.  The grid is ordered tile
// fastest so that concurrently running CTAs read the same part of the kernel spectrum (L2 hits).
template <int BZ> struct ZCfg {
#ifndef MVD_Z_EPT512
#define MVD_Z_EPT512 16
#endif
    // elements per thread: 512 = 16 x 16 x 2 (two exchanges), 256 = 16 x 16, 128 = 16 x 8, 64 = 8 x 8 (one exchange)
    static constexpr int EPT = BZ >= 512 ? MVD_Z_EPT512 : (BZ >= 128 ? 16 : 8);
    static constexpr int TZ = BZ / EPT;
#ifndef MVD_Z_PW
#define MVD_Z_PW 16
#endif
    static constexpr int PW = MVD_Z_PW;  // 16: 128 contiguous bytes per z index (32 / 64 / 128 / 256 bytes: 16.4 / 14.8 / 11.1 / 12.8 ms)
    static constexpr int THREADS = PW * TZ;
    // exchange buffer | twiddles
    static constexpr size_t SMEM = (size_t)PW * BZ * sizeof(float2) + (size_t)BZ * sizeof(float2);
#ifndef MVD_Z_MINCTAS512
#define MVD_Z_MINCTAS512 (1024 / THREADS)
#endif
    // shorter lines: 3 x 256 threads at <= 85 registers (measured faster than 4 x 256 at 64)
    static constexpr int MIN_CTAS_ = BZ >= 512 ? MVD_Z_MINCTAS512 : (THREADS >= 256 ? 3 : 4);
    static constexpr int MIN_CTAS = MIN_CTAS_ < 1 ? 1 : MIN_CTAS_;
};

// Work item -> (position block, tile): z_items.cuh.  Two round-1 experiments lived here and were removed in round 2 once the
// TMA kernels superseded them (both measured slower): per-thread cp.async staging of the next item
// (MVD_Z_ASYNC, 14.0 - 17.4 vs 12.5 ms per iteration) and a register software pipeline (MVD_Z_PIPE, 12.5 vs 11.1 ms).
template <int BZ, int MODE>
__global__ void __launch_bounds__(ZCfg<BZ>::THREADS, ZCfg<BZ>::MIN_CTAS) z_pass_kernel(float2* __restrict__ scratch, size_t tile_stride,
                                                                                     const float2* __restrict__ kspec,
                                                                                     const Tile* __restrict__ tiles, int plane_elems,
                                                                                     float scale, int n_tiles, int n_items, Twiddles tw) {
    using C = ZCfg<BZ>;
    constexpr int EPT = C::EPT, TZ = C::TZ, PW = C::PW, NTH = C::THREADS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* xch = reinterpret_cast<float2*>(smem_raw);
    float2* twz = xch + (size_t)PW * BZ;
    const int tid = threadIdx.x;
    const int p = tid % PW, t = tid / PW;
    for (int i = tid; i < BZ; i += NTH) twz[i] = __ldg(&tw.t[ilog2<BZ>()][i]);

#pragma unroll 1
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    // work items = (position block, tile), tile fastest: the CTAs running at the same time read the
    // same part of the kernel spectrum (L2 hits)
    int pblk, ti;
    decode_item(item, n_tiles, pblk, ti);
    const int pos = pblk * PW + p;
    const bool valid = pos < plane_elems;
    float2* __restrict__ base = scratch + (size_t)ti * tile_stride + (valid ? pos : 0);

    // Element offsets are recomputed in every phase from a laundered copy of the plane size: the
    // loads, the spectrum reads and the stores use the same EPT addresses, and common-subexpression
    // elimination would otherwise keep all of them (2 registers each) alive across both transforms.
    float2 v[EPT];
    {
        const unsigned pl = (unsigned)plane_elems;
#pragma unroll
        for (int m = 0; m < EPT; ++m) v[m] = base[(unsigned)(t + m * TZ) * pl];
    }
    const float2* __restrict__ kb = kspec + (valid ? pos : 0);
    __syncthreads();  // twiddle table ready; the previous item is done with the exchange buffer

    float2* __restrict__ xp = xch + p;
    auto put = [&](int idx, float2 val) { xp[idx * PW] = val; };
    auto get = [&](int idx) { return xp[idx * PW]; };
    auto sync = [] { __syncthreads(); };
    line_fft<BZ, EPT, false>(v, t, twz, put, get, sync);

    if constexpr (MODE == 0) {
        if (valid) {
            const unsigned pl = (unsigned)opaque(plane_elems);
#pragma unroll
            for (int m = 0; m < EPT; ++m) base[(unsigned)(t + m * TZ) * pl] = make_float2(v[m].x * scale, v[m].y * scale);
        }
    } else {
        // spectrum values in batches of KB: all EPT at once would need 2 * EPT more registers than the
        // transform itself and spill (the compiler hoists every load above the first multiplication)
#ifndef MVD_Z_KB
#define MVD_Z_KB 4
#endif
#ifndef MVD_Z_KB_SHORT
#define MVD_Z_KB_SHORT 16  // shorter lines run at <= 85 registers: no batching needed (measured)
#endif
        constexpr int KB_ = BZ >= 512 ? MVD_Z_KB : MVD_Z_KB_SHORT;
        constexpr int KB = EPT >= KB_ ? KB_ : EPT;
#pragma unroll
        for (int h = 0; h < EPT; h += KB) {
            float2 kk[KB];
            const unsigned pl = (unsigned)opaque(plane_elems);
#pragma unroll
            for (int i = 0; i < KB; ++i) kk[i] = __ldg(kb + (unsigned)(t + (h + i) * TZ) * pl);
#pragma unroll
            for (int i = 0; i < KB; ++i) v[h + i] = cmul(v[h + i], kk[i]);
            asm volatile("" ::: "memory");
        }
        if constexpr (TZ > 1) __syncthreads();
        line_fft<BZ, EPT, true>(v, t, twz, put, get, sync);
        const Tile tile = tiles[ti];
        const int zlo = tile.e[2] - tile.o[2];
        const int zhi = zlo + tile.n[2];
        if (valid) {
            const unsigned pl = (unsigned)opaque(plane_elems);
#pragma unroll
            for (int m = 0; m < EPT; ++m) {
                const int z = t + m * TZ;
                if (z >= zlo && z < zhi) base[(unsigned)z * pl] = v[m];
            }
        }
    }
    }  // items
}

template <int BX, int BY, int EPI, bool PUSH>
__global__ void __launch_bounds__((InvThreads<BX, BY>::NT), (InvCfg<BX, BY>::MIN_CTAS < InvThreads<BX, BY>::MIN_CTAS_WANT ? InvCfg<BX, BY>::MIN_CTAS : InvThreads<BX, BY>::MIN_CTAS_WANT)) inv_xy_kernel(const float2* __restrict__ scratch, size_t tile_stride,
                                                                           const Tile* __restrict__ tiles, EpilogueArgs ep, int chunks,
                                                                           int n_items, Twiddles tw) {
    using C = InvCfg<BX, BY>;
    using L = PlaneLayout<BX>;
    constexpr int NT = C::NT;
    constexpr int H = C::H, P = L::P, TX = C::TX, RPR = C::RPR, TY = C::TY, CPR = C::CPR;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* plane = reinterpret_cast<float2*>(smem_raw);
    float2* twr = plane + (size_t)BY * P;
    float2* twu = twr + H;
    __shared__ double s_sum[NT / 32];
    __shared__ float s_max[NT / 32];
    __shared__ PeerTarget s_peer[PUSH ? kMaxPeers : 1];  // PUSH: neighbours whose ghost layers intersect the current tile
    __shared__ int s_npeer;

    const int tid = threadIdx.x;
    if constexpr (BX == 128) {
        row128_fill_tables(twr, twu, tw.t[6], tw.t[7], true, tid, NT);
    } else {
        for (int i = tid; i < H; i += NT) {
            twr[i] = __ldg(&tw.t[ilog2<H>()][i]);
            twu[i] = __ldg(&tw.t[ilog2<BX>()][i]);
        }
    }
    const int rt = tid % TX;
    const int rrow = tid / TX;
    const int cc = tid % CPR;
    const int cu = tid / CPR;
    const ConstTw256 twc{8 - ilog2<BY>()};
    const bool use_tik = ep.lambda > 0.0;
    const TikhonovConst tik = {ep.tik_hi, ep.tik_lo, ep.lambda};
    double lsum = 0.0;
    float lmax = 0.f;
    __syncthreads();
    // work items = (tile, chunk of XY_ZPC effective planes), chunk fastest (see fwd_xy_kernel)
#pragma unroll 1
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int ti = item / chunks;
    const Tile tile = tiles[ti];
    const int zfirst = (item - ti * chunks) * XY_ZPC;
    if (zfirst >= tile.n[2]) continue;
    const int lx = tile.e[0] - tile.o[0], ly = tile.e[1] - tile.o[1];
    const int n0 = tile.n[0], n1 = tile.n[1];
    if constexpr (PUSH) {
        // read by the row phase behind the barrier that ends the column phase; the previous item's
        // rows are behind the barrier that ended its last plane
        if (tid < 32) {
            bool hit = false;
            PeerTarget t;
            if (tid < ep.n_peers) {
                t = ep.peers[tid];
                hit = true;
#pragma unroll
                for (int d = 0; d < 3; ++d) hit = hit && tile.e[d] < t.hi[d] && tile.e[d] + tile.n[d] > t.lo[d];
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) s_peer[__popc(m & ((1u << tid) - 1u))] = t;
            if (tid == 0) s_npeer = __popc(m);
        }
    }
    // x pairs owned by this thread: x = 2*(rt + m*TX) - lx, +1; classify once (lx is a multiple of 4)
    unsigned full = 0, part = 0;
#pragma unroll
    for (int m = 0; m < 8; ++m) {
        const int xl = 2 * (rt + m * TX) - lx;
        if (xl >= 0 && xl + 1 < n0) full |= 1u << m;
        else if (xl + 1 == 0 || xl == n0 - 1) part |= 1u << m;
    }

#pragma unroll 1
    for (int zi = 0; zi < XY_ZPC; ++zi) {
        const int zr = zfirst + zi;
        if (zr >= tile.n[2]) break;
        const int zt = tile.e[2] - tile.o[2] + zr;  // plane inside the tile
        const int zv = tile.e[2] + zr;              // plane inside the volume
        const float2* __restrict__ in = scratch + (size_t)ti * tile_stride + (size_t)zt * ((H + 1) * BY);

        if constexpr (BX == 128 && (MVD_PREFETCH & 6) != 0) {
            // pull this plane's epilogue operands (2) and the next plane's spectrum (4) towards L2
            if (MVD_PREFETCH & 2)
            for (int yl = tid >> 2; yl < n1; yl += NT / 4) {
                const int q = tid & 3;
                const size_t ob = ((size_t)(zv + ep.out_off[2]) * ep.out_dim[1] + (tile.e[1] + yl + ep.out_off[1])) * (size_t)ep.out_dim[0] +
                                  tile.e[0] + ep.out_off[0] + 32 * q;
                const size_t ab = ((size_t)zv * ep.aux_dim[1] + (tile.e[1] + yl)) * (size_t)ep.aux_dim[0] + tile.e[0] + 32 * q;
                if (32 * q < n0) {
                    if constexpr (EPI == EPI_QUOTIENT) prefetch_l2(ep.img + ab);
                    if constexpr (EPI == EPI_UPDATE) {
                        prefetch_l2(ep.out + ob);
                        if (ep.weight) prefetch_l2(ep.weight + ab);
                    }
                }
            }
            if ((MVD_PREFETCH & 4) && zi + 1 < XY_ZPC && zr + 1 < tile.n[2]) {
                const float2* __restrict__ nx = in + (H + 1) * BY;
                for (int i = tid; i < (H + 1) * BY / 16; i += NT) prefetch_l2(nx + 16 * i);
            }
        }
        // ---- phase 1: inverse columns; column 0 re-packs DC + i*Nyquist
#pragma unroll 1
        for (int cr = 0; cr < C::CROUNDS; ++cr) {
            const int col = cr * CPR + cc;
            const bool valid = C::CFULL || col < H;
            float2 v[16];
#pragma unroll
            for (int m = 0; m < 16; ++m) v[m] = valid ? __ldg(in + (cu + m * TY) * H + col) : make_float2(0.f, 0.f);
            if (col == 0) {
#pragma unroll
                for (int m = 0; m < 16; ++m) {
                    const float2 b = __ldg(in + H * BY + cu + m * TY);
                    v[m] = make_float2(v[m].x - b.y, v[m].y + b.x);
                }
            }
            float2* __restrict__ cp = plane + L::slot(valid ? col : 0);
            auto put = [&](int idx, float2 val) { if (valid) cp[idx * P] = val; };
            auto get = [&](int idx) { return cp[idx * P]; };
            auto sync = [] { __syncthreads(); };
            line_fft<BY, 16, true>(v, cu, twc, put, get, sync);
            if constexpr (TY > 1) __syncthreads();
#pragma unroll
            for (int m = 0; m < 16; ++m) put(cu + m * TY, v[m]);
        }
        __syncthreads();

        // ---- phase 2: inverse rows (C2R through the packed half-length transform) + epilogue.
        // Only the rows of the effective region are transformed; the others would be discarded.
        if constexpr (BX == 128) {
            inv_rows128<EPI, PUSH, NT>(plane, twr, twu, tile, zv, ep, s_peer, s_npeer, tid, use_tik, tik, lsum, lmax);
        } else {
#pragma unroll 1
        for (int yl0 = 0; yl0 < n1; yl0 += RPR) {
            const int yl = yl0 + rrow;
            // rows of the effective region come in multiples of 4, a warp holds 32/TX of them
            const unsigned wmask = __ballot_sync(0xffffffffu, yl < n1);
            if (yl >= n1) continue;
            const int row = ly + yl;
            int pxlo = 0, pxhi = 0;
            const unsigned pm = PUSH && s_npeer ? push_row_mask(s_peer, s_npeer, tile.e[1] + yl, zv, pxlo, pxhi) : 0u;
            float2* __restrict__ rp = plane + row * P;
            float2 v[8];
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                const int k = rt + m * TX;
                const float2 xk = rp[L::slot(k)];
                const float2 xp = cconj(rp[L::slot((H - k) & (H - 1))]);
                const float2 s = cadd(xk, xp), d = csub(xk, xp);
                const float2 wd = cmulc(d, twu[k]);
                v[m] = make_float2(s.x - wd.y, s.y + wd.x);  // s + i*conj(w)*d
                if (m == 0 && rt == 0) v[0] = make_float2(xk.x + xk.y, xk.x - xk.y);
            }
            __syncwarp(wmask);
            auto put = [&](int idx, float2 val) { rp[L::slot(idx)] = val; };
            auto get = [&](int idx) { return rp[L::slot(idx)]; };
            auto sync = [&] { __syncwarp(wmask); };
            line_fft<H, 8, true>(v, rt, twr, put, get, sync);

            const size_t rowbase = ((size_t)(zv + ep.out_off[2]) * ep.out_dim[1] + (tile.e[1] + yl + ep.out_off[1])) * (size_t)ep.out_dim[0] +
                                   tile.e[0] + ep.out_off[0];
            const size_t auxbase = ((size_t)zv * ep.aux_dim[1] + (tile.e[1] + yl)) * (size_t)ep.aux_dim[0] + tile.e[0];
            // a pair is 8-byte aligned whenever the row start is
            const bool vec = ((rowbase | auxbase) & 1) == 0;
            float* __restrict__ outp = ep.out + rowbase + 2 * rt - lx;
            const float* __restrict__ imgp = ep.img + auxbase + 2 * rt - lx;
            const float* __restrict__ wp = ep.weight + auxbase + 2 * rt - lx;
            unsigned rest = part;
            if (vec) {
                float2 a[8], b[8];
                // issue every load before the first use
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    a[m] = make_float2(0.f, 0.f);
                    b[m] = make_float2(ep.weight_const, ep.weight_const);
                    if (full >> m & 1) {
                        if constexpr (EPI == EPI_QUOTIENT) a[m] = __ldg(reinterpret_cast<const float2*>(imgp + 2 * m * TX));
                        if constexpr (EPI == EPI_UPDATE) {
                            a[m] = *reinterpret_cast<const float2*>(outp + 2 * m * TX);
                            if (ep.weight) b[m] = __ldg(reinterpret_cast<const float2*>(wp + 2 * m * TX));
                        }
                    }
                }
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    if (full >> m & 1) {
                        const float c0 = v[m].x + ep.add_const, c1 = v[m].y + ep.add_const;
                        float2 r;
                        if constexpr (EPI == EPI_STORE) r = make_float2(c0, c1);
                        else if constexpr (EPI == EPI_QUOTIENT) r = make_float2(quotient_value(a[m].x, c0), quotient_value(a[m].y, c1));
                        else {
                            r.x = next_psi_value(a[m].x, c0, b[m].x, use_tik, tik, ep.min_value);
                            r.y = next_psi_value(a[m].y, c1, b[m].y, use_tik, tik, ep.min_value);
                            const float ch0 = fabsf(__fsub_rn(r.x, a[m].x)), ch1 = fabsf(__fsub_rn(r.y, a[m].y));
                            lsum += (double)(ch0 + ch1);
                            lmax = fmaxf(lmax, fmaxf(ch0, ch1));  // fmaxf drops NaN
                        }
                        *reinterpret_cast<float2*>(outp + 2 * m * TX) = r;
                        if constexpr (PUSH) {
                            const int px = tile.e[0] + 2 * (rt + m * TX) - lx;
                            if (pm && px + 2 > pxlo && px < pxhi) push_run<2>(s_peer, pm, px, tile.e[1] + yl, zv, make_float4(r.x, r.y, 0.f, 0.f));
                        }
                    }
                }
            } else {
                rest |= full;
            }
            // scalar path: pairs straddling the region border, or unaligned rows
            if (rest) {
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    if (!(rest >> m & 1)) continue;
                    const int xl = 2 * (rt + m * TX) - lx;
                    const float c[2] = {v[m].x + ep.add_const, v[m].y + ep.add_const};
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int x = xl + q;
                        if (x < 0 || x >= n0) continue;
                        const size_t idx = rowbase + x, aidx = auxbase + x;
                        float nv;
                        if constexpr (EPI == EPI_STORE) {
                            nv = c[q];
                        } else if constexpr (EPI == EPI_QUOTIENT) {
                            nv = quotient_value(__ldg(ep.img + aidx), c[q]);
                        } else {
                            const float last = ep.out[idx];
                            const float w = ep.weight ? __ldg(ep.weight + aidx) : ep.weight_const;
                            nv = next_psi_value(last, c[q], w, use_tik, tik, ep.min_value);
                            const float ch = fabsf(__fsub_rn(nv, last));
                            lsum += (double)ch;
                            lmax = fmaxf(lmax, ch);
                        }
                        ep.out[idx] = nv;
                        if constexpr (PUSH) {
                            if (pm && tile.e[0] + x >= pxlo && tile.e[0] + x < pxhi) push_run<1>(s_peer, pm, tile.e[0] + x, tile.e[1] + yl, zv, make_float4(nv, 0.f, 0.f, 0.f));
                        }
                    }
                }
            }
        }
        }
        __syncthreads();  // rows are done with the plane before the next plane's columns scatter into it
        if constexpr (PUSH && BX == 128) {
            if (s_npeer) {  // the same for every thread of the CTA (written behind a barrier): interior tiles pay nothing
                push_plane128<NT>(plane, tile, zv, s_peer, s_npeer, tid);
                __syncthreads();
            }
        }
    }
    }  // items
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

// ---------------------------------------------------------------------------------------------
// host launchers
// grid caps (0 = one CTA per work item): with several batch streams the kernels run as persistent
// grids of a fraction of the SMs' capacity, so CTAs of different kernels share an SM
int g_xy_grid_cap = 0;
int g_z_grid_cap = 0;
int g_num_sms = 148;  // set by configure_conv_kernels
void set_grid_caps(int xy_cap, int z_cap) { g_xy_grid_cap = xy_cap; g_z_grid_cap = z_cap; }

bool xy_size_supported(int b) { return b == 32 || b == 64 || b == 128; }
bool y_size_supported(int b) { return xy_size_supported(b) || b == 256; }
bool xy_pair_supported(int bx, int by) { return y_size_supported(bx) && y_size_supported(by) && !(bx == 256 && by == 256); }
bool z_size_supported(int b) { return b == 16 || b == 32 || b == 64 || b == 128 || b == 256 || b == 512; }
size_t plane_elems(int bx, int by) { return (size_t)(bx / 2 + 1) * by; }
size_t tile_scratch_elems(int bx, int by, int bz) { return plane_elems(bx, by) * bz; }

static inline Twiddles to_dev(const TwiddleTables& t) {
    Twiddles d;
    for (int i = 0; i < 11; ++i) d.t[i] = t.t[i];
    return d;
}

#define MVD_XY_DISPATCH(bx, by, F)                                                            \
    do {                                                                                      \
        if (bx == 32 && by == 32) { F(32, 32); } else if (bx == 32 && by == 64) { F(32, 64); } \
        else if (bx == 32 && by == 128) { F(32, 128); } else if (bx == 64 && by == 32) { F(64, 32); } \
        else if (bx == 64 && by == 64) { F(64, 64); } else if (bx == 64 && by == 128) { F(64, 128); } \
        else if (bx == 128 && by == 32) { F(128, 32); } else if (bx == 128 && by == 64) { F(128, 64); } \
        else if (bx == 128 && by == 128) { F(128, 128); } else if (bx == 32 && by == 256) { F(32, 256); }  \
        else if (bx == 64 && by == 256) { F(64, 256); } else if (bx == 128 && by == 256) { F(128, 256); } \
        else if (bx == 256 && by == 32) { F(256, 32); } else if (bx == 256 && by == 64) { F(256, 64); }   \
        else if (bx == 256 && by == 128) { F(256, 128); }                                                 \
        else return cudaErrorInvalidValue;      \
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
    {
        int dev = 0, sms = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0)
            g_num_sms = sms;
    }
    // column twiddles in the constant bank
    {
        float2 h[256];
        const double two_pi = 6.283185307179586476925286766559;
        for (int i = 0; i < 256; ++i) {
            const double a = -two_pi * (double)i / 256.0;
            h[i] = make_float2((float)cos(a), (float)sin(a));
        }
        e = cudaMemcpyToSymbol(c_tw256, h, sizeof(h));
    }
#define CFG_XY(BX, BY)                                                                                                           \
    {                                                                                                                            \
        constexpr int sm = (int)XYCfg<BX, BY>::SMEM_INV;                                                                         \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(fwd_xy_kernel<BX, BY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)XYCfg<BX, BY>::SMEM_FWD); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_STORE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_QUOTIENT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_UPDATE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_QUOTIENT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_xy_kernel<BX, BY, EPI_UPDATE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm); \
    }
    CFG_XY(32, 32) CFG_XY(32, 64) CFG_XY(32, 128) CFG_XY(64, 32) CFG_XY(64, 64) CFG_XY(64, 128) CFG_XY(128, 32) CFG_XY(128, 64) CFG_XY(128, 128)
    CFG_XY(32, 256) CFG_XY(64, 256) CFG_XY(128, 256) CFG_XY(256, 32) CFG_XY(256, 64) CFG_XY(256, 128)
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
    const int n_items = ((bz + XY_ZPC - 1) / XY_ZPC) * n_tiles;
    const int grid = g_xy_grid_cap > 0 ? std::min(n_items, g_xy_grid_cap) : n_items;
#define F(BX, BY) fwd_xy_kernel<BX, BY><<<grid, XYCfg<BX, BY>::NT, XYCfg<BX, BY>::SMEM_FWD, st>>>(src, d_tiles, scratch, stride, bz, n_items, to_dev(tw))
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
        const int n_pblk = (pe + ZCfg<BZ>::PW - 1) / ZCfg<BZ>::PW;                                                     \
        const int n_items = n_tiles * ((n_pblk + MVD_Z_PGROUP - 1) / MVD_Z_PGROUP * MVD_Z_PGROUP);                     \
        const int cap = g_z_grid_cap;                                                                                  \
        const int grid = cap > 0 ? std::min(n_items, cap) : n_items;                                                   \
        if (mode == 0) z_pass_kernel<BZ, 0><<<grid, ZCfg<BZ>::THREADS, ZCfg<BZ>::SMEM, st>>>(scratch, stride, kspec, d_tiles, pe, scale, n_tiles, n_items, to_dev(tw)); \
        else z_pass_kernel<BZ, 1><<<grid, ZCfg<BZ>::THREADS, ZCfg<BZ>::SMEM, st>>>(scratch, stride, kspec, d_tiles, pe, scale, n_tiles, n_items, to_dev(tw)); \
    }
    MVD_Z_DISPATCH(bz, F);
#undef F
    return cudaGetLastError();
}

cudaError_t launch_inv_xy(int bx, int by, int bz, int epilogue, const float2* scratch, const Tile* d_tiles, int n_tiles, int max_nz,
                          const EpilogueArgs& ep, const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0 || max_nz <= 0) return cudaSuccess;
    if (ep.n_peers > kMaxPeers) return cudaErrorInvalidValue;
    const size_t stride = tile_scratch_elems(bx, by, bz);
    const int chunks = (max_nz + XY_ZPC - 1) / XY_ZPC;
    const int n_items = chunks * n_tiles;
    const int grid = g_xy_grid_cap > 0 ? std::min(n_items, g_xy_grid_cap) : n_items;
#define F(BX, BY)                                                                                                        \
    {                                                                                                                    \
        constexpr size_t sm = XYCfg<BX, BY>::SMEM_INV;                                                                       \
        const bool push = ep.peers != nullptr && ep.n_peers > 0;                                                         \
        if (epilogue == EPI_STORE) inv_xy_kernel<BX, BY, EPI_STORE, false><<<grid, InvCfg<BX, BY>::NT, sm, st>>>(scratch, stride, d_tiles, ep, chunks, n_items, to_dev(tw)); \
        else if (epilogue == EPI_QUOTIENT && !push) inv_xy_kernel<BX, BY, EPI_QUOTIENT, false><<<grid, InvCfg<BX, BY>::NT, sm, st>>>(scratch, stride, d_tiles, ep, chunks, n_items, to_dev(tw)); \
        else if (epilogue == EPI_QUOTIENT) inv_xy_kernel<BX, BY, EPI_QUOTIENT, true><<<grid, InvCfg<BX, BY>::NT, sm, st>>>(scratch, stride, d_tiles, ep, chunks, n_items, to_dev(tw)); \
        else if (!push) inv_xy_kernel<BX, BY, EPI_UPDATE, false><<<grid, InvCfg<BX, BY>::NT, sm, st>>>(scratch, stride, d_tiles, ep, chunks, n_items, to_dev(tw)); \
        else inv_xy_kernel<BX, BY, EPI_UPDATE, true><<<grid, InvCfg<BX, BY>::NT, sm, st>>>(scratch, stride, d_tiles, ep, chunks, n_items, to_dev(tw)); \
    }
    MVD_XY_DISPATCH(bx, by, F);
#undef F
    return cudaGetLastError();
}

}  // namespace mvd
