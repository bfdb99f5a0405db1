// Row phase of the inverse xy pass for 128-wide windows with the fused per-voxel epilogue, shared by the
// register-loading kernel (conv_kernels.cu) and the ring kernel fed by bulk copies (inv_ring_kernels.cu).
#pragma once
#include "conv_types.h"
#include "device_common.cuh"
#include "fft_device.cuh"
#include "row128.cuh"

namespace mvd {

// Fused ghost-layer exchange: a run of N freshly computed output voxels starting at brick
// coordinate (x0, y, z) is also stored into the ghost layers of every neighbouring brick that
// needs it (PeerTarget, conv_types.h) -- peer memory over NVLink, written by the same kernel
// that produced the values.  `peers` is the tile's candidate list in shared memory and `pm` the
// subset whose y/z range contains the row (push_row_mask), so interior tiles never get here.
// also returns the x range [xlo, xhi) covered by the matching neighbours, so that runs outside it skip the call
__device__ __forceinline__ unsigned push_row_mask(const PeerTarget* peers, int n, int y, int z, int& xlo, int& xhi) {
    unsigned m = 0;
    xlo = 0x7fffffff;
    xhi = -0x7fffffff;
    for (int i = 0; i < n; ++i) {
        const PeerTarget& t = peers[i];
        if (y >= t.lo[1] && y < t.hi[1] && z >= t.lo[2] && z < t.hi[2]) {
            m |= 1u << i;
            xlo = min(xlo, t.lo[0]);
            xhi = max(xhi, t.hi[0]);
        }
    }
    return m;
}

template <int N>
__device__ __noinline__ void push_run(const PeerTarget* __restrict__ peers, unsigned pm, int x0, int y, int z, float4 r) {
    const float rv[4] = {r.x, r.y, r.z, r.w};
    for (; pm; pm &= pm - 1) {
        const PeerTarget& t = peers[__ffs(pm) - 1];
        if (x0 + N <= t.lo[0] || x0 >= t.hi[0]) continue;
        const long long idx = ((long long)(z + t.shift[2]) * t.dim[1] + (y + t.shift[1])) * (long long)t.dim[0] + (x0 + t.shift[0]);
        float* __restrict__ p = t.base + idx;
        const bool inside = x0 >= t.lo[0] && x0 + N <= t.hi[0];
        if (N == 4 && inside && (idx & 3) == 0) {
            *reinterpret_cast<float4*>(p) = r;
        } else if (N == 2 && inside && (idx & 1) == 0) {
            *reinterpret_cast<float2*>(p) = make_float2(r.x, r.y);
        } else {
#pragma unroll
            for (int i = 0; i < N; ++i)
                if (x0 + i >= t.lo[0] && x0 + i < t.hi[0]) p[i] = rv[i];
        }
    }
}


// Inverse rows (C2R through the packed half-length transform) of the effective region of one plane + epilogue.
// `plane`: BY rows of 64 packed complex columns, pitch 68 (PlaneLayout<128>); NTG threads (lt = 0 .. NTG-1) share the
// plane, four per row.  Only the rows of the effective region are transformed; the others would be discarded.
template <int EPI, bool PUSH, int NTG>
__device__ __forceinline__ void inv_rows128(float2* plane, const float2* twr, const float2* twu, const Tile& tile, int zv, const EpilogueArgs& ep,
                                            const PeerTarget* s_peer, int npeer, int lt, bool use_tik, const TikhonovConst& tik, double& lsum,
                                            float& lmax) {
    const int lx = tile.e[0] - tile.o[0], ly = tile.e[1] - tile.o[1];
    const int n0 = tile.n[0], n1 = tile.n[1];
    const int q = lt & 3;
#pragma unroll 1
    for (int yl0 = 0; yl0 < n1; yl0 += NTG / 4) {
        const int yl = yl0 + (lt >> 2);
        // rows of the effective region come in multiples of 4, a warp holds 8 of them
        const unsigned wmask = __ballot_sync(0xffffffffu, yl < n1);
        if (yl >= n1) continue;
        float2 e[8], f[8];
        float rsum = 0.f;  // |change| of this thread's <= 32 voxels of the row, added to the double sum once
        // PUSH: the results also go into the row's (by then consumed) shared-memory storage, float index lx + x;
        // push_plane128 copies the neighbours' parts out once the whole plane is done
        float* const rowf = reinterpret_cast<float*>(plane + (ly + yl) * 68);
        const size_t rowbase = ((size_t)(zv + ep.out_off[2]) * ep.out_dim[1] + (tile.e[1] + yl + ep.out_off[1])) * (size_t)ep.out_dim[0] +
                               tile.e[0] + ep.out_off[0];
        const size_t auxbase = ((size_t)zv * ep.aux_dim[1] + (tile.e[1] + yl)) * (size_t)ep.aux_dim[0] + tile.e[0];
        const bool vec = ((rowbase | auxbase) & 3) == 0 && ((reinterpret_cast<uintptr_t>(ep.out) | reinterpret_cast<uintptr_t>(ep.img) |
                                                              reinterpret_cast<uintptr_t>(ep.weight)) & 15) == 0;
        const int xq = 4 * q - lx;  // first sample of the thread's quad m = 0, in region coordinates
        unsigned full4 = 0, part4 = 0;
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const int xl = xq + 16 * m;
            if (xl >= 0 && xl + 3 < n0) full4 |= 1u << m;
            else if (xl + 3 >= 0 && xl < n0) part4 |= 1u << m;
        }
        float* __restrict__ outp = ep.out + rowbase + xq;
        const float* __restrict__ imgp = ep.img + auxbase + xq;
        const float* __restrict__ wp = ep.weight + auxbase + xq;
        // epilogue operands of quads h .. h+3 (the first half is requested before the row transform,
        // the second half before the first half is consumed: two batches of loads in flight)
        auto load_half = [&](int h, float4 (&a)[4], float4 (&b)[4]) {
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
                const int m = h + i4;
                a[i4] = make_float4(0.f, 0.f, 0.f, 0.f);
                b[i4] = make_float4(ep.weight_const, ep.weight_const, ep.weight_const, ep.weight_const);
                if (vec && (full4 >> m & 1)) {
                    if constexpr (EPI == EPI_QUOTIENT) a[i4] = __ldg(reinterpret_cast<const float4*>(imgp + 16 * m));
                    if constexpr (EPI == EPI_UPDATE) {
                        a[i4] = *reinterpret_cast<const float4*>(outp + 16 * m);
                        if (ep.weight) b[i4] = __ldg(reinterpret_cast<const float4*>(wp + 16 * m));
                    }
                }
            }
        };
        auto finish_half = [&](int h, const float4 (&a)[4], const float4 (&b)[4]) {
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
                const int m = h + i4;
                if (full4 >> m & 1) {
                    const float c[4] = {e[m].x + ep.add_const, e[m].y + ep.add_const, f[m].x + ep.add_const, f[m].y + ep.add_const};
                    const float av[4] = {a[i4].x, a[i4].y, a[i4].z, a[i4].w};
                    const float bv[4] = {b[i4].x, b[i4].y, b[i4].z, b[i4].w};
                    float r[4];
                    if constexpr (EPI == EPI_UPDATE) {
                        next_psi_quad(av, c, bv, use_tik, tik, ep.min_value, r);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float ch = fabsf(__fsub_rn(r[i], av[i]));
                            rsum += ch;
                            lmax = fmaxf(lmax, ch);  // fmaxf drops NaN
                        }
                    } else if constexpr (EPI == EPI_QUOTIENT) {
                        quotient_quad(av, c, r);
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i) r[i] = c[i];
                    }
                    *reinterpret_cast<float4*>(outp + 16 * m) = make_float4(r[0], r[1], r[2], r[3]);
                    if (PUSH && npeer) *reinterpret_cast<float4*>(rowf + 4 * q + 16 * m) = make_float4(r[0], r[1], r[2], r[3]);
                }
            }
        };
        float4 a0[4], b0[4];
        load_half(0, a0, b0);
        row128_inverse(e, f, plane + (ly + yl) * 68, twr, twu, q, wmask);
        if (PUSH && npeer) __syncwarp(wmask);  // every thread of the row has gathered its samples: the row's storage is free
        // (e[m].x, e[m].y, f[m].x, f[m].y) = samples x = 16 m + 4 q .. + 3 of the window
        if (vec) {
            float4 a1[4], b1[4];
            load_half(4, a1, b1);
            finish_half(0, a0, b0);
            finish_half(4, a1, b1);
        } else {
            part4 |= full4;
        }
        if (part4) {  // quads straddling the region border, or unaligned rows
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                if (!(part4 >> m & 1)) continue;
                const float c[4] = {e[m].x + ep.add_const, e[m].y + ep.add_const, f[m].x + ep.add_const, f[m].y + ep.add_const};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int x = xq + 16 * m + i;
                    if (x < 0 || x >= n0) continue;
                    const size_t idx = rowbase + x, aidx = auxbase + x;
                    float nv;
                    if constexpr (EPI == EPI_STORE) {
                        nv = c[i];
                    } else if constexpr (EPI == EPI_QUOTIENT) {
                        nv = quotient_value(__ldg(ep.img + aidx), c[i]);
                    } else {
                        const float last = ep.out[idx];
                        const float w = ep.weight ? __ldg(ep.weight + aidx) : ep.weight_const;
                        nv = next_psi_value(last, c[i], w, use_tik, tik, ep.min_value);
                        const float ch = fabsf(__fsub_rn(nv, last));
                        rsum += ch;
                        lmax = fmaxf(lmax, ch);
                    }
                    ep.out[idx] = nv;
                    if (PUSH && npeer) rowf[lx + x] = nv;
                }
            }
        }
        if constexpr (EPI == EPI_UPDATE) lsum += (double)rsum;
    }
}

// Staged ghost-layer push of one plane (PUSH kernels), called by all NTG threads behind the barrier that ends the plane's
// row phase: every neighbour's sub-rectangle of the tile's effective region is copied from the plane's rows (where
// inv_rows128 left the results) into the neighbour's ghost layers -- peer memory over NVLink -- with consecutive lanes on
// consecutive x.  The epilogue itself carries no per-quad tests or calls any more (they made boundary tiles ~70 % slower
// although only ~16 % of their voxels are pushed).
template <int NTG>
__device__ __forceinline__ void push_plane128(const float2* plane, const Tile& tile, int zv, const PeerTarget* s_peer, int npeer, int lt) {
    const int lx = tile.e[0] - tile.o[0], ly = tile.e[1] - tile.o[1];
    const float* const pf = reinterpret_cast<const float*>(plane);
    for (int i = 0; i < npeer; ++i) {
        const PeerTarget& t = s_peer[i];
        if (zv < t.lo[2] || zv >= t.hi[2]) continue;
        const int xa = max(tile.e[0], t.lo[0]), xb = min(tile.e[0] + tile.n[0], t.hi[0]);
        const int ya = max(tile.e[1], t.lo[1]), yb = min(tile.e[1] + tile.n[1], t.hi[1]);
        const int w = xb - xa, h = yb - ya;
        if (w <= 0 || h <= 0) continue;
        float* __restrict__ dst = t.base + ((long long)(zv + t.shift[2]) * t.dim[1] + (ya + t.shift[1])) * (long long)t.dim[0] + (xa + t.shift[0]);
        const float* __restrict__ src = pf + (ly + ya - tile.e[1]) * 136 + (lx + xa - tile.e[0]);
        const int pitch = t.dim[0];
        if (((w | (lx + xa - tile.e[0])) & 3) == 0 && ((pitch | (xa + t.shift[0])) & 3) == 0 && (reinterpret_cast<uintptr_t>(t.base) & 15) == 0) {
            const int w4 = w >> 2;
            for (int j = lt; j < w4 * h; j += NTG) {
                const int yy = j / w4, xx = (j - yy * w4) << 2;
                *reinterpret_cast<float4*>(dst + (long long)yy * pitch + xx) = *reinterpret_cast<const float4*>(src + yy * 136 + xx);
            }
        } else {
            for (int j = lt; j < w * h; j += NTG) {
                const int yy = j / w, xx = j - yy * w;
                dst[(long long)yy * pitch + xx] = src[yy * 136 + xx];
            }
        }
    }
}

}  // namespace mvd
