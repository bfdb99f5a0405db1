// z pass of the tiled FFT convolution with warp-local lines (sm_100a), 512-point windows.
//
// The register-loading z pass (conv_kernels.cu) and its TMA variant (z_tma_kernels.cu) spread a 512-point line over
// 32 threads of 16 different warps, so that lanes run across plane positions and global accesses coalesce; every
// exchange of the line FFT is then a barrier of 512 threads, ten per work item, and an SM holds only two such groups.
// Here a line lives in ONE warp (lane t holds z = t + 32 m) and the transposition lanes-across-positions ->
// lanes-across-z happens in shared memory behind the TMA unit:
//
//   * a work item (16 plane positions x 512 planes, rows of 128 B that lie plane_elems * 8 bytes apart) is landed
//     by two cp.async.bulk.tensor loads with the 128-byte swizzle, warp w takes position w out of the landing buffer
//     (4 instead of 2 shared-memory wavefronts per access thanks to the swizzle, 32 without it);
//   * the three passes of the line FFT (16 x 16 x 2) exchange through a warp-private, padded line buffer behind
//     __syncwarp -- no CTA or group barrier anywhere; the inter-pass twiddles of the second pass depend on the lane
//     only and stay in registers;
//   * the kernel spectrum is kept TRANSPOSED ([position][z]) so that a warp reads its line of it coalesced, and a CTA
//     works through all tiles of one position block before it moves on: the line's 16 spectrum values per lane stay in
//     registers across those tiles (no spectrum traffic per item at all);
//   * results go to a staging buffer in the swizzled box layout and leave with two bulk tensor stores; the landing
//     buffer is refilled as soon as the last warp has taken its line (counter in shared memory, the last arriver
//     issues the copy), so the next item's load overlaps the whole transform.
//
// Replaces the z part of J/fftMethods/FFTConvolution.java:470-551 (C2C along z, multiply :607-614, inverse).
#include "conv_kernels.h"
#include "fft_device.cuh"
#include "tma.cuh"

#include <algorithm>

namespace mvd {

// BZ = 512: 16 elements per lane, passes 16 x 16 x 2; BZ = 256: 8 elements per lane, passes 8 x 8 x 4
template <int BZ> struct ZW {
    static constexpr int EPT = BZ / 32;                // elements per lane
    static constexpr int R2 = BZ / (EPT * EPT);        // radix of the third pass (2 or 4)
    static constexpr int NB = EPT / R2;                // butterflies per lane in the third pass
    static constexpr int PW = 16, THREADS = PW * 32;
    static constexpr int XP = BZ + BZ / EPT;           // padded private line: one pad element per EPT
    static constexpr unsigned SLOT_BYTES = PW * BZ * sizeof(float2);  // 65536 / 32768
    static constexpr int NBOX = BZ / 256;              // a box holds at most 256 rows
    static constexpr size_t OFF_L = 0, OFF_ST = SLOT_BYTES, OFF_X = 2 * (size_t)SLOT_BYTES;
    static constexpr size_t OFF_TW3 = OFF_X + (size_t)PW * XP * sizeof(float2);
    static constexpr size_t OFF_BAR = OFF_TW3 + 256 * sizeof(float2);
    static constexpr size_t SMEM = OFF_BAR + 64;
    static constexpr int MIN_CTAS = BZ >= 512 ? 1 : 2;  // 200 KB / 101 KB of shared memory
    static_assert(BZ == 512 || BZ == 256, "warp-local lines: 512 or 256 points");
};
// Tiles a CTA does per position block before it takes the next unit (the line's kernel-spectrum values stay in registers
// across them): chosen per launch so that the most loaded CTA has the least work.
static int choose_tiles_per_unit(int n_tiles, int n_pblk, int grid) {
    int best_tc = 1;
    double best = 1e300;
    for (int tc = 1; tc <= 16 && tc <= n_tiles; ++tc) {
        const int chunks = (n_tiles + tc - 1) / tc;
        const long long units = (long long)chunks * n_pblk;
        const long long rounds = (units + grid - 1) / grid;
        const double cost = (double)rounds * (tc + 0.5);  // + 0.5: a unit reloads its spectrum values (half an item's traffic)
        if (cost < best) { best = cost; best_tc = tc; }
    }
    return best_tc;
}

// byte offset of element (z, p) inside a landing / staging buffer written or read by the TMA unit with the 128-byte
// swizzle: rows of 128 B (16 positions), the 16-byte chunk index is XORed with the row index modulo 8
__device__ __forceinline__ unsigned zw_swz(int z, int p) {
    return (unsigned)z * 128u + ((((unsigned)p >> 1) ^ ((unsigned)z & 7u)) << 4) + ((unsigned)p & 1u) * 8u;
}

// Passes 2 and 3 of the line FFT inside a warp, after the leading radix-EPT butterfly: lane t holds the
// butterfly's outputs in v on entry and X[t + 32 m] on return.  Same passes and arithmetic as
// line_fft<BZ, EPT, INV> (fft_device.cuh); xw = the warp's padded line buffer (pad(i) = i + i / EPT: every access
// pattern below is conflict free for 8-byte accesses).
template <int BZ, bool INV>
__device__ __forceinline__ void warp_fft_tail(float2 (&v)[ZW<BZ>::EPT], int t, float2* __restrict__ xw, const float2 (&tw2)[ZW<BZ>::EPT - 1],
                                              const float2* __restrict__ tw3) {
    constexpr int EPT = ZW<BZ>::EPT, R2 = ZW<BZ>::R2, NB = ZW<BZ>::NB;
    constexpr int LG = EPT == 16 ? 4 : 3;
    {   // pass 1 output r of lane t -> position EPT t + r; pad(EPT t + r) = (EPT + 1) t + r
        float2* p = xw + (EPT + 1) * t;
#pragma unroll
        for (int r = 0; r < EPT; ++r) p[r] = v[r];
    }
    __syncwarp();
    const float2* g = xw + t + (t >> LG);  // pad(t + 32 m) = t + (t >> LG) + (32 + 32 / EPT) m
    constexpr int GS = 32 + 32 / EPT;
#pragma unroll
    for (int m = 0; m < EPT; ++m) v[m] = g[GS * m];
    __syncwarp();
    // pass 2: j = t, k = t & (EPT - 1), twiddles w^(k r BZ / EPT^2), radix EPT; output r -> position EPT (t - k) + k + EPT r
#pragma unroll
    for (int r = 1; r < EPT; ++r) v[r] = ctw<INV>(v[r], tw2[r - 1]);
    dftR<EPT, INV>(v);
    {
        float2* p = xw + (EPT + 1) * EPT * (t >> LG) + (t & (EPT - 1));  // pad(...) = (EPT + 1)(t - k) + k + (EPT + 1) r
#pragma unroll
        for (int r = 0; r < EPT; ++r) p[(EPT + 1) * r] = v[r];
    }
    __syncwarp();
#pragma unroll
    for (int m = 0; m < EPT; ++m) v[m] = g[GS * m];
    __syncwarp();  // the buffer may be overwritten by the next transform
    // pass 3: radix R2 on (v[b], v[b + NB], ...), j = t + 32 b, twiddles w^(j r)
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        float2 x[R2];
#pragma unroll
        for (int r = 0; r < R2; ++r) x[r] = v[b + r * NB];
#pragma unroll
        for (int r = 1; r < R2; ++r) x[r] = ctw<INV>(x[r], tw3[(t + 32 * b) * r]);
        dftR<R2, INV>(x);
#pragma unroll
        for (int r = 0; r < R2; ++r) v[b + r * NB] = x[r];
    }
}

template <int BZ>
__global__ void __launch_bounds__(ZW<BZ>::THREADS, ZW<BZ>::MIN_CTAS)
    z_warp_kernel(const __grid_constant__ CUtensorMap map_ld, const __grid_constant__ CUtensorMap map_st, const float2* __restrict__ kspec_t,
                  int n_tiles, int n_pblk, int n_chunks, int TC, Twiddles tw) {
    using C = ZW<BZ>;
    constexpr int EPT = C::EPT, PW = C::PW, THREADS = C::THREADS, XP = C::XP;
    constexpr unsigned SLOT_BYTES = C::SLOT_BYTES;
    constexpr size_t OFF_L = C::OFF_L, OFF_ST = C::OFF_ST, OFF_X = C::OFF_X, OFF_TW3 = C::OFF_TW3, OFF_BAR = C::OFF_BAR;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* L = smem_raw + OFF_L;
    unsigned char* ST = smem_raw + OFF_ST;
    float2* tw3 = reinterpret_cast<float2*>(smem_raw + OFF_TW3);
    uint64_t* full_l = reinterpret_cast<uint64_t*>(smem_raw + OFF_BAR);
    uint64_t* st_free = full_l + 1;
    int* cnt_l = reinterpret_cast<int*>(full_l + 2);
    int* cnt_st = cnt_l + 1;

    const int tid = threadIdx.x;
    const int w = tid >> 5, t = tid & 31;  // warp = plane position inside the block, lane = z residue
    float2* xw = reinterpret_cast<float2*>(smem_raw + OFF_X) + (size_t)w * XP;
    for (int i = tid; i < 256; i += THREADS) tw3[i] = __ldg(&tw.t[ilog2<BZ>()][i]);
    float2 tw2[EPT - 1];
#pragma unroll
    for (int r = 1; r < EPT; ++r) tw2[r - 1] = __ldg(&tw.t[ilog2<BZ>()][(BZ / (EPT * EPT)) * (t & (EPT - 1)) * r]);
    if (tid == 0) {
        tma::mbar_init(full_l, 1);
        tma::mbar_init(st_free, 1);
        *cnt_l = 0;
        *cnt_st = 0;
        tma::fence_barrier_init();
    }
    __syncthreads();

    // units = (tile chunk, position block), position block fastest: CTAs running side by side read neighbouring
    // position blocks of the same tiles.  A CTA takes units blockIdx.x, + gridDim.x, ... and inside a unit the tiles
    // of the chunk in order.
    const int n_units = n_pblk * n_chunks;
    struct Item { int unit, tile; };
    auto first_tile = [&](int unit) { return (unit / n_pblk) * TC; };
    auto last_tile = [&](int unit) { return min(n_tiles, (unit / n_pblk + 1) * TC); };
    auto advance = [&](Item& it) {
        if (++it.tile >= last_tile(it.unit)) {
            it.unit += gridDim.x;
            it.tile = first_tile(it.unit);
        }
    };
    auto issue_load = [&](const Item& it) {  // one thread
        const int pblk = it.unit % n_pblk;
        tma::mbar_arrive_expect_tx(full_l, SLOT_BYTES);
#pragma unroll
        for (int b = 0; b < C::NBOX; ++b) tma::load_3d(L + (size_t)b * 256 * 128, &map_ld, full_l, 2 * PW * pblk, 256 * b, it.tile);
    };
    Item it{(int)blockIdx.x, 0};
    it.tile = first_tile(it.unit);
    if (tid == 0 && it.unit < n_units) issue_load(it);

    // this lane's EPT elements sit 32 rows = 4096 bytes apart in a swizzled buffer
    const unsigned lane_off = zw_swz(t, w);
    int cur_unit = -1;
    float2 kreg[EPT];
    unsigned seq = 0;
#pragma unroll 1
    for (; it.unit < n_units; advance(it), ++seq) {
        if (it.unit != cur_unit) {  // the line's part of the kernel spectrum (1/n folded in), coalesced from the transposed copy
            cur_unit = it.unit;
            const float2* __restrict__ kp = kspec_t + ((size_t)(it.unit % n_pblk) * PW + w) * BZ + t;
#pragma unroll
            for (int m = 0; m < EPT; ++m) kreg[m] = __ldg(kp + 32 * m);
        }
        tma::mbar_wait(full_l, seq & 1u);
        float2 v[EPT];
#pragma unroll
        for (int m = 0; m < EPT; ++m) v[m] = *reinterpret_cast<const float2*>(L + lane_off + 4096u * m);
        // forward transform; its first butterflies consume every element, so the landing buffer is released right behind them
        dftR<EPT, false>(v);
        __syncwarp();
        if (t == 0) {
            __threadfence_block();
            if (atomicAdd(cnt_l, 1) == PW - 1) {  // the last warp to take its line refills the buffer with the next item
                *cnt_l = 0;
                Item nx = it;
                advance(nx);
                if (nx.unit < n_units) {
                    tma::fence_proxy_async();
                    issue_load(nx);
                }
            }
        }
        warp_fft_tail<BZ, false>(v, t, xw, tw2, tw3);
#pragma unroll
        for (int m = 0; m < EPT; ++m) v[m] = cmul(v[m], kreg[m]);
        dftR<EPT, true>(v);
        warp_fft_tail<BZ, true>(v, t, xw, tw2, tw3);
        // results -> staging buffer (free once the previous item's store has read it), swizzled box layout
        tma::mbar_wait(st_free, (seq & 1u) ^ 1u);
#pragma unroll
        for (int m = 0; m < EPT; ++m) *reinterpret_cast<float2*>(ST + lane_off + 4096u * m) = v[m];
        tma::fence_proxy_async();
        __syncwarp();
        if (t == 0) {
            __threadfence_block();
            if (atomicAdd(cnt_st, 1) == PW - 1) {  // the last warp to deliver its line sends the item off
                *cnt_st = 0;
                const int pblk = it.unit % n_pblk;
                tma::fence_proxy_async();
#pragma unroll
                for (int b = 0; b < C::NBOX; ++b) tma::store_3d(&map_st, ST + (size_t)b * 256 * 128, 2 * PW * pblk, 256 * b, it.tile);
                tma::store_commit();
                tma::store_wait_read();
                tma::mbar_arrive(st_free);
            }
        }
    }
    if (t == 0) tma::store_wait_all();
}

// ---------------------------------------------------------------------------------------------------------------------
// Second generation: a line lives in a PART of a warp (BZ / 32 lanes, 32 elements per lane), so that the line FFT has two
// passes (32 x 16 for 512 points, 32 x 8 for 256) and ONE exchange instead of three passes and two exchanges, and a warp
// holds 32 / (BZ / 32) neighbouring plane positions.  The first generation is bound by the shared-memory pipe (l1tex 74 %
// busy in ncu: 4 wavefronts per 256 bytes at the swizzled buffers, two exchanges per transform); here
//   * every half-warp (the unit of an 8-byte shared-memory access) holds 8 different z residues of TWO neighbouring
//     positions, i.e. both halves of 8 different 16-byte chunks: 16 bank pairs, no conflict, at the landing rows, at the
//     result rows and -- through the row permutation xrow() -- at the exchange;
//   * an item (16 positions x BZ planes) lands in one of THREE buffers and stays there: a warp takes its lines out of its own
//     16-byte column of the buffer, exchanges inside that column, leaves the results in it, and the last warp to finish
//     sends the buffer off (bulk tensor store) and refills it with the item three ahead -- two loads are in flight per
//     CTA while one item is computed, and warps never wait for each other except through the refill;
//   * half as many threads per item (8 warps for 512 points), so there is room for 255 registers per thread: the 32 spectrum
//     values per lane and all inter-pass twiddles stay in registers.
template <int BZ> struct ZH {
    static constexpr int EPT = 32;
    static constexpr int LPL = BZ / EPT;               // lanes per line (16 / 8)
    static constexpr int LPW = 32 / LPL;               // lines (plane positions) per warp (2 / 4)
    static constexpr int R1 = LPL;                     // radix of the second pass
    static constexpr int NB = EPT / R1;                // butterflies per lane in the second pass (2 / 4)
    static constexpr int PW = 16, NWARP = PW / LPW, THREADS = NWARP * 32;
    static constexpr int NBUF = 3;
    static constexpr unsigned SLOT_BYTES = PW * BZ * sizeof(float2);  // 65536 / 32768
    static constexpr int NBOX = BZ / 256;
    static constexpr size_t OFF_BAR = (size_t)NBUF * SLOT_BYTES;
    static constexpr size_t SMEM = OFF_BAR + 64;
    static constexpr int MIN_CTAS = BZ >= 512 ? 1 : 2;  // 192 KB / 96 KB of shared memory
    static_assert(BZ == 512 || BZ == 256, "part-warp lines: 512 or 256 points");
};

// Exchange row of line element i: the low three bits are XORed with bits 5..7, so that the 8 lanes of a half-warp that
// belong to one line touch 8 different rows modulo 8 (= 8 different swizzled chunks) when they write 32 l + r (fixed r) and
// when they read l + LPL m (fixed m).
__host__ __device__ constexpr int xrow(int i) { return i ^ ((i >> 5) & 7); }

// One line FFT of BZ points on LPL lanes: lane l of the line holds x[l + LPL m] in v[m] on entry and X[l + LPL m] on return
// (the passes of line_fft<BZ, 32, INV>, fft_device.cuh: radix 32 with NS = 1, then radix LPL with NS = 32 and twiddles
// w^(j r), j = l + LPL b).  buf + xo[c]: the byte address, in the item's swizzled buffer, of row (c ^ l & 7) of this
// lane's column; put_base / get_base: the lane's part of the row offset (32 l rows / l & 8 rows).
template <int BZ, bool INV>
__device__ __forceinline__ void part_warp_fft(float2 (&v)[32], unsigned char* __restrict__ buf, const unsigned (&xo)[8], unsigned put_base,
                                              unsigned get_base, const float2 (&tw)[ZH<BZ>::NB][ZH<BZ>::R1 - 1]) {
    constexpr int LPL = ZH<BZ>::LPL, R1 = ZH<BZ>::R1, NB = ZH<BZ>::NB;
    dftR<32, INV>(v);
    // output r of lane l -> element 32 l + r -> row 32 l + (r ^ (l & 7)) = 32 l + 8 (r >> 3) + ((r & 7) ^ (l & 7))
#pragma unroll
    for (int r = 0; r < 32; ++r) *reinterpret_cast<float2*>(buf + put_base + xo[r & 7] + 1024u * (r >> 3)) = v[r];
    __syncwarp();
    // element i = l + LPL m: i >> 5 = (LPL m) >> 5 =: c (l < LPL <= 16 never carries), row = (i & ~7) + ((l & 7) ^ (c & 7))
#pragma unroll
    for (int m = 0; m < 32; ++m) {
        const int c = (LPL * m) >> 5;
        v[m] = *reinterpret_cast<const float2*>(buf + get_base + xo[c & 7] + 128u * ((LPL * m) & ~7));
    }
    __syncwarp();  // the column may be overwritten by the next exchange or by the results
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        float2 x[R1];
#pragma unroll
        for (int r = 0; r < R1; ++r) x[r] = v[b + r * NB];
#pragma unroll
        for (int r = 1; r < R1; ++r) x[r] = ctw<INV>(x[r], tw[b][r - 1]);
        dftR<R1, INV>(x);
#pragma unroll
        for (int r = 0; r < R1; ++r) v[b + r * NB] = x[r];
    }
}

template <int BZ>
__global__ void __launch_bounds__(ZH<BZ>::THREADS, ZH<BZ>::MIN_CTAS)
    z_part_kernel(const __grid_constant__ CUtensorMap map_ld, const __grid_constant__ CUtensorMap map_st, const float2* __restrict__ kspec_t,
                  int n_tiles, int n_pblk, int n_chunks, int TC, Twiddles tw_tab) {
    using C = ZH<BZ>;
    constexpr int LPL = C::LPL, LPW = C::LPW, PW = C::PW, NWARP = C::NWARP, R1 = C::R1, NB = C::NB, NBUF = C::NBUF;
    constexpr unsigned SLOT_BYTES = C::SLOT_BYTES;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::OFF_BAR);  // [NBUF]
    int* cnt = reinterpret_cast<int*>(full + NBUF);                       // [NBUF]

    const int tid = threadIdx.x;
    const int w = tid >> 5, t = tid & 31;
    // lane -> (line h inside the warp, lane l inside the line); a half-warp holds l & 7 = 0..7 of two lines
    const int h = LPL == 16 ? (t >> 3) & 1 : t >> 3;
    const int l = LPL == 16 ? (t & 7) | ((t >> 4) << 3) : t & 7;
    const int p = LPW * w + h;           // plane position inside the block of 16
    float2 tw[NB][R1 - 1];
#pragma unroll
    for (int b = 0; b < NB; ++b)
#pragma unroll
        for (int r = 1; r < R1; ++r) tw[b][r - 1] = __ldg(&tw_tab.t[ilog2<BZ>()][(l + LPL * b) * r]);
    if (tid == 0) {
#pragma unroll
        for (int b = 0; b < NBUF; ++b) {
            tma::mbar_init(full + b, 1);
            cnt[b] = 0;
        }
        tma::fence_barrier_init();
    }
    __syncthreads();

    const int n_units = n_pblk * n_chunks;
    struct Item { int unit, tile; };
    auto first_tile = [&](int unit) { return (unit / n_pblk) * TC; };
    auto last_tile = [&](int unit) { return min(n_tiles, (unit / n_pblk + 1) * TC); };
    auto advance = [&](Item& it) {
        if (++it.tile >= last_tile(it.unit)) {
            it.unit += gridDim.x;
            it.tile = first_tile(it.unit);
        }
    };
    auto issue_load = [&](const Item& it, int b) {  // one thread
        const int pblk = it.unit % n_pblk;
        unsigned char* dst = smem_raw + (size_t)b * SLOT_BYTES;
        tma::mbar_arrive_expect_tx(full + b, SLOT_BYTES);
#pragma unroll
        for (int x = 0; x < C::NBOX; ++x) tma::load_3d(dst + (size_t)x * 256 * 128, &map_ld, full + b, 2 * PW * pblk, 256 * x, it.tile);
    };
    Item it{(int)blockIdx.x, 0};
    it.tile = first_tile(it.unit);
    if (tid == 0) {
        Item nx = it;
        for (int b = 0; b < NBUF && nx.unit < n_units; ++b, advance(nx)) issue_load(nx, b);
    }

    // rows of this lane's column: z = l + LPL m for the item's data, xrow() rows for the exchange; z & 7 = l & 7
    const unsigned pc = (unsigned)p >> 1, hb = ((unsigned)p & 1u) * 8u, l7 = (unsigned)l & 7u;
    const unsigned lane_off = zw_swz(l, p);
    constexpr unsigned MSTEP = LPL * 128u;
    unsigned xo[8];
#pragma unroll
    for (unsigned c = 0; c < 8; ++c) xo[c] = ((c ^ l7) << 7) + (((pc ^ l7 ^ c) & 7u) << 4) + hb;
    const unsigned put_base = (unsigned)l * 4096u, get_base = ((unsigned)l & 8u) * 128u;
    int cur_unit = -1;
    float2 kreg[32];
    unsigned seq = 0;
    int b = 0;
#pragma unroll 1
    for (; it.unit < n_units; advance(it), ++seq, b = b == NBUF - 1 ? 0 : b + 1) {
        if (it.unit != cur_unit) {  // the line's kernel-spectrum values (1/n folded in) from the transposed copy
            cur_unit = it.unit;
            const float2* __restrict__ kp = kspec_t + ((size_t)(it.unit % n_pblk) * PW + p) * BZ + l;
#pragma unroll
            for (int m = 0; m < 32; ++m) kreg[m] = __ldg(kp + LPL * m);
        }
        unsigned char* buf = smem_raw + (size_t)b * SLOT_BYTES;
        tma::mbar_wait(full + b, (seq / NBUF) & 1u);
        float2 v[32];
#pragma unroll
        for (int m = 0; m < 32; ++m) v[m] = *reinterpret_cast<const float2*>(buf + lane_off + MSTEP * m);
        __syncwarp();  // every lane has its elements before the column becomes exchange space
        part_warp_fft<BZ, false>(v, buf, xo, put_base, get_base, tw);
#pragma unroll
        for (int m = 0; m < 32; ++m) v[m] = cmul(v[m], kreg[m]);
        part_warp_fft<BZ, true>(v, buf, xo, put_base, get_base, tw);
#pragma unroll
        for (int m = 0; m < 32; ++m) *reinterpret_cast<float2*>(buf + lane_off + MSTEP * m) = v[m];
        tma::fence_proxy_async();
        __syncwarp();
        if (t == 0) {
            __threadfence_block();
            int old;
            asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(tma::smem_u32(cnt + b)) : "memory");
            if (old == NWARP - 1) {  // the last warp to deliver its lines sends the item off and refills the buffer
                cnt[b] = 0;
                __threadfence_block();
                const int pblk = it.unit % n_pblk;
                tma::fence_proxy_async();
#pragma unroll
                for (int x = 0; x < C::NBOX; ++x) tma::store_3d(&map_st, buf + (size_t)x * 256 * 128, 2 * PW * pblk, 256 * x, it.tile);
                tma::store_commit();
                Item nx = it;
#pragma unroll
                for (int a = 0; a < NBUF; ++a) advance(nx);
                if (nx.unit < n_units) {
                    tma::store_wait_read();
                    issue_load(nx, b);
                }
            }
        }
    }
    if (t == 0) tma::store_wait_all();
}

// [bz][pe] -> [pe][bz]
__global__ void transpose_spectrum_kernel(const float2* __restrict__ in, float2* __restrict__ out, int pe, int bz) {
    __shared__ float2 tile[32][33];
    const int p0 = blockIdx.x * 32, z0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int z = z0 + i, p = p0 + threadIdx.x;
        if (z < bz && p < pe) tile[i][threadIdx.x] = in[(size_t)z * pe + p];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int p = p0 + i, z = z0 + threadIdx.x;
        if (z < bz && p < pe) out[(size_t)p * bz + z] = tile[threadIdx.x][i];
    }
}

bool z_warp_supported(int bz) { return bz == 512 || bz == 256; }

cudaError_t configure_z_warp_kernels() {
    cudaError_t e = cudaFuncSetAttribute(z_warp_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZW<512>::SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(z_warp_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZW<256>::SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(z_part_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZH<512>::SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(z_part_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZH<256>::SMEM);
    return e;
}

cudaError_t launch_transpose_spectrum(const float2* spec, float2* spec_t, int pe, int bz, cudaStream_t st) {
    dim3 grid((pe + 31) / 32, (bz + 31) / 32), block(32, 8);
    transpose_spectrum_kernel<<<grid, block, 0, st>>>(spec, spec_t, pe, bz);
    return cudaGetLastError();
}

static inline Twiddles to_dev(const TwiddleTables& t) {
    Twiddles d;
    for (int i = 0; i < 11; ++i) d.t[i] = t.t[i];
    return d;
}

// cudaErrorNotSupported: the caller falls back to the register-loading z pass
cudaError_t launch_z_pass_warp(int bx, int by, int bz, float2* scratch, const float2* kspec_t, int n_tiles, int zlo, int max_nz, int num_sms,
                               const TwiddleTables& tw, cudaStream_t st, int generation) {
    if (n_tiles <= 0) return cudaSuccess;
    if (!z_warp_supported(bz) || !kspec_t) return cudaErrorNotSupported;
    const unsigned long long pe = plane_elems(bx, by);
    if (pe % 16 != 0) return cudaErrorNotSupported;
    const int zext = std::min(bz, zlo + max_nz);
    CUtensorMap ld, stm;
    const unsigned long long strides[2] = {pe * 8ull, pe * 8ull * (unsigned long long)bz};
    const unsigned long long dims_ld[3] = {2ull * pe, (unsigned long long)bz, (unsigned long long)n_tiles};
    const unsigned long long dims_st[3] = {2ull * pe, (unsigned long long)zext, (unsigned long long)n_tiles};
    const unsigned box[3] = {32u, 256u, 1u};
    if (!encode_tensor_map_f32_swizzle128(&ld, scratch, 3, dims_ld, strides, box) ||
        !encode_tensor_map_f32_swizzle128(&stm, scratch, 3, dims_st, strides, box))
        return cudaErrorNotSupported;
    const int n_pblk = (int)(pe / 16);
    const int cap = num_sms * (bz == 512 ? 1 : 2);  // MIN_CTAS of both generations
    const int tc = choose_tiles_per_unit(n_tiles, n_pblk, cap);
    const int n_chunks = (n_tiles + tc - 1) / tc;
    const int grid = std::max(1, std::min(cap, n_pblk * n_chunks));
    if (generation >= 2) {
        if (bz == 512) z_part_kernel<512><<<grid, ZH<512>::THREADS, ZH<512>::SMEM, st>>>(ld, stm, kspec_t, n_tiles, n_pblk, n_chunks, tc, to_dev(tw));
        else z_part_kernel<256><<<grid, ZH<256>::THREADS, ZH<256>::SMEM, st>>>(ld, stm, kspec_t, n_tiles, n_pblk, n_chunks, tc, to_dev(tw));
    } else if (bz == 512) z_warp_kernel<512><<<grid, ZW<512>::THREADS, ZW<512>::SMEM, st>>>(ld, stm, kspec_t, n_tiles, n_pblk, n_chunks, tc, to_dev(tw));
    else z_warp_kernel<256><<<grid, ZW<256>::THREADS, ZW<256>::SMEM, st>>>(ld, stm, kspec_t, n_tiles, n_pblk, n_chunks, tc, to_dev(tw));
    return cudaGetLastError();
}

}  // namespace mvd
