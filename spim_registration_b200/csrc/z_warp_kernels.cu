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

namespace zw {
constexpr int BZ = 512, PW = 16, THREADS = PW * 32;
constexpr int XP = BZ + BZ / 16;                           // padded private line: one pad element per 16
constexpr unsigned SLOT_BYTES = PW * BZ * sizeof(float2);  // 65536
constexpr size_t OFF_L = 0, OFF_ST = SLOT_BYTES, OFF_X = 2 * (size_t)SLOT_BYTES;
constexpr size_t OFF_TW3 = OFF_X + (size_t)PW * XP * sizeof(float2);
constexpr size_t OFF_BAR = OFF_TW3 + 256 * sizeof(float2);
constexpr size_t SMEM = OFF_BAR + 64;
constexpr int TC = 11;  // tiles a CTA does per position block before it takes the next unit
}  // namespace zw

// byte offset of element (z, p) inside a landing / staging buffer written or read by the TMA unit with the 128-byte
// swizzle: rows of 128 B (16 positions), the 16-byte chunk index is XORed with the row index modulo 8
__device__ __forceinline__ unsigned zw_swz(int z, int p) {
    return (unsigned)z * 128u + ((((unsigned)p >> 1) ^ ((unsigned)z & 7u)) << 4) + ((unsigned)p & 1u) * 8u;
}

// Passes 2 and 3 of the 512-point line FFT inside a warp, after the leading radix-16 butterfly: lane t holds the
// butterfly's outputs in v on entry and X[t + 32 m] on return.  Same passes and arithmetic as
// line_fft<512, 16, INV> (fft_device.cuh); xw = the warp's padded line buffer.
template <bool INV>
__device__ __forceinline__ void warp_fft512_tail(float2 (&v)[16], int t, float2* __restrict__ xw, const float2 (&tw2)[15],
                                                 const float2* __restrict__ tw3) {
    {   // pass 1 output r of lane t -> position 16 t + r; pad(16 t + r) = 17 t + r
        float2* p = xw + 17 * t;
#pragma unroll
        for (int r = 0; r < 16; ++r) p[r] = v[r];
    }
    __syncwarp();
    const float2* g = xw + t + (t >> 4);  // pad(t + 32 m) = t + (t >> 4) + 34 m
#pragma unroll
    for (int m = 0; m < 16; ++m) v[m] = g[34 * m];
    __syncwarp();
    // pass 2: j = t, k = t & 15, twiddles w512^(2 k r), radix 16; output r -> position 256 (t >> 4) + k + 16 r
#pragma unroll
    for (int r = 1; r < 16; ++r) v[r] = ctw<INV>(v[r], tw2[r - 1]);
    dft16<INV>(v);
    {
        float2* p = xw + 272 * (t >> 4) + (t & 15);  // pad(...) = 272 (t >> 4) + k + 17 r
#pragma unroll
        for (int r = 0; r < 16; ++r) p[17 * r] = v[r];
    }
    __syncwarp();
#pragma unroll
    for (int m = 0; m < 16; ++m) v[m] = g[34 * m];
    __syncwarp();  // the buffer may be overwritten by the next transform
    // pass 3: radix 2 on (v[b], v[b + 8]), j = t + 32 b, twiddle w512^j
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const float2 x1 = ctw<INV>(v[b + 8], tw3[t + 32 * b]);
        const float2 x0 = v[b];
        v[b] = cadd(x0, x1);
        v[b + 8] = csub(x0, x1);
    }
}

__global__ void __launch_bounds__(zw::THREADS, 1)
    z_warp_kernel(const __grid_constant__ CUtensorMap map_ld, const __grid_constant__ CUtensorMap map_st, const float2* __restrict__ kspec_t,
                  int n_tiles, int n_pblk, int n_chunks, Twiddles tw) {
    using namespace zw;
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
    for (int i = tid; i < 256; i += THREADS) tw3[i] = __ldg(&tw.t[9][i]);
    float2 tw2[15];
#pragma unroll
    for (int r = 1; r < 16; ++r) tw2[r - 1] = __ldg(&tw.t[9][2 * (t & 15) * r]);
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
        tma::load_3d(L, &map_ld, full_l, 2 * PW * pblk, 0, it.tile);
        tma::load_3d(L + SLOT_BYTES / 2, &map_ld, full_l, 2 * PW * pblk, 256, it.tile);
    };
    Item it{(int)blockIdx.x, 0};
    it.tile = first_tile(it.unit);
    if (tid == 0 && it.unit < n_units) issue_load(it);

    // this lane's 16 elements sit 32 rows = 4096 bytes apart in a swizzled buffer
    const unsigned lane_off = zw_swz(t, w);
    int cur_unit = -1;
    float2 kreg[16];
    unsigned seq = 0;
#pragma unroll 1
    for (; it.unit < n_units; advance(it), ++seq) {
        if (it.unit != cur_unit) {  // the line's part of the kernel spectrum (1/n folded in), coalesced from the transposed copy
            cur_unit = it.unit;
            const float2* __restrict__ kp = kspec_t + ((size_t)(it.unit % n_pblk) * PW + w) * BZ + t;
#pragma unroll
            for (int m = 0; m < 16; ++m) kreg[m] = __ldg(kp + 32 * m);
        }
        tma::mbar_wait(full_l, seq & 1u);
        float2 v[16];
#pragma unroll
        for (int m = 0; m < 16; ++m) v[m] = *reinterpret_cast<const float2*>(L + lane_off + 4096u * m);
        // forward transform; its first butterflies consume every element, so the landing buffer is released right behind them
        dft16<false>(v);
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
        warp_fft512_tail<false>(v, t, xw, tw2, tw3);
#pragma unroll
        for (int m = 0; m < 16; ++m) v[m] = cmul(v[m], kreg[m]);
        dft16<true>(v);
        warp_fft512_tail<true>(v, t, xw, tw2, tw3);
        // results -> staging buffer (free once the previous item's store has read it), swizzled box layout
        tma::mbar_wait(st_free, (seq & 1u) ^ 1u);
#pragma unroll
        for (int m = 0; m < 16; ++m) *reinterpret_cast<float2*>(ST + lane_off + 4096u * m) = v[m];
        tma::fence_proxy_async();
        __syncwarp();
        if (t == 0) {
            __threadfence_block();
            if (atomicAdd(cnt_st, 1) == PW - 1) {  // the last warp to deliver its line sends the item off
                *cnt_st = 0;
                const int pblk = it.unit % n_pblk;
                tma::fence_proxy_async();
                tma::store_3d(&map_st, ST, 2 * PW * pblk, 0, it.tile);
                tma::store_3d(&map_st, ST + SLOT_BYTES / 2, 2 * PW * pblk, 256, it.tile);
                tma::store_commit();
                tma::store_wait_read();
                tma::mbar_arrive(st_free);
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

bool z_warp_supported(int bz) { return bz == zw::BZ; }

cudaError_t configure_z_warp_kernels() {
    return cudaFuncSetAttribute(z_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)zw::SMEM);
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
                               const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0) return cudaSuccess;
    if (!z_warp_supported(bz) || !kspec_t) return cudaErrorNotSupported;
    const unsigned long long pe = plane_elems(bx, by);
    if (pe % zw::PW != 0) return cudaErrorNotSupported;
    const int zext = std::min(bz, zlo + max_nz);
    CUtensorMap ld, stm;
    const unsigned long long strides[2] = {pe * 8ull, pe * 8ull * (unsigned long long)bz};
    const unsigned long long dims_ld[3] = {2ull * pe, (unsigned long long)bz, (unsigned long long)n_tiles};
    const unsigned long long dims_st[3] = {2ull * pe, (unsigned long long)zext, (unsigned long long)n_tiles};
    const unsigned box[3] = {32u, 256u, 1u};
    if (!encode_tensor_map_f32_swizzle128(&ld, scratch, 3, dims_ld, strides, box) ||
        !encode_tensor_map_f32_swizzle128(&stm, scratch, 3, dims_st, strides, box))
        return cudaErrorNotSupported;
    const int n_pblk = (int)(pe / zw::PW);
    const int n_chunks = (n_tiles + zw::TC - 1) / zw::TC;
    const int grid = std::max(1, std::min(num_sms, n_pblk * n_chunks));
    z_warp_kernel<<<grid, zw::THREADS, zw::SMEM, st>>>(ld, stm, kspec_t, n_tiles, n_pblk, n_chunks, to_dev(tw));
    return cudaGetLastError();
}

}  // namespace mvd
