// Inverse xy pass for 128 x 128 windows fed by bulk copies into a shared-memory ring (sm_100a).
//
// Same arithmetic as inv_xy_kernel<128, 128, EPI, PUSH> (conv_kernels.cu): inverse column transforms, inverse rows
// (C2R) of the effective region and the fused per-voxel epilogue (store / quotient / multiplicative update,
// MVDeconvolution.java:473-525, 582-703).  What changes is how a spectrum plane reaches the transform and how the
// threads of a CTA are coupled:
//
//   * a plane (128 rows x 64 columns + the Nyquist column, 66.5 KB contiguous in the scratch) is fetched by ONE thread with
//     two cp.async.bulk.tensor copies (UTMALDG) straight into the padded plane layout of a shared-memory slot and
//     announced on an mbarrier: the scratch is described as rows of 64 complex values and the box is 68 wide, so the
//     out-of-range columns are zero-filled and the rows land with the conflict-free pitch of 68 -- no thread holds a
//     load in a register or computes a scratch address (129 row-wise cp.async.bulk copies per plane were tried first:
//     the issuing warp then reaches every group barrier late, 15.9 vs 12.0 ms per four convolutions);
//   * the CTA holds S = 3 slots and G = 2 independent groups of 256 threads (named barriers): a group transforms a
//     plane in place in its slot while the next planes are in flight, and refills the slot the moment its row phase has
//     consumed it;
//   * column transforms are warp-local: the eight threads of a column sit in one warp (four columns per warp), the
//     exchange between the two passes is ordered by __syncwarp and uses a row swizzle that keeps every shared-memory
//     access conflict free, so the warps of a group run out of step.  A plane costs two group barriers (columns -> rows,
//     rows -> slot free) where the register-loading kernel needs six CTA barriers.
//
// CTAs are persistent and own a contiguous run of (tile, plane) pairs.
#include "conv_kernels.h"
#include "fft_device.cuh"
#include "inv_rows128.cuh"
#include "row128.cuh"
#include "tma.cuh"

#include <algorithm>

namespace mvd {

namespace ring {
constexpr int BY = 128, H = 64, P = 68;
constexpr int GT = 256, G = 2, S = 3, THREADS = G * GT;
constexpr int SLOT_ELEMS = BY * P + BY;                      // padded plane + Nyquist column
constexpr unsigned SLOT_BYTES = SLOT_ELEMS * sizeof(float2);  // 70656
constexpr unsigned PLANE_BYTES = SLOT_BYTES;                  // bytes that arrive per plane, zero-filled pad columns included
constexpr int ROWS_PER_PLANE = BY + BY / H;  // scratch rows of 64 complex per plane: 128 + the Nyquist column (2 rows)
// slots | row tables (64 + 32) | w128 table (128) | peers (G x kMaxPeers) | mbarriers
constexpr size_t OFF_TW = (size_t)S * SLOT_BYTES;
constexpr size_t OFF_W128 = OFF_TW + 96 * sizeof(float2);
constexpr size_t OFF_PEER = OFF_W128 + 128 * sizeof(float2);
constexpr size_t OFF_BAR = OFF_PEER + (size_t)G * kMaxPeers * sizeof(PeerTarget);
constexpr size_t SMEM = OFF_BAR + 64;
}  // namespace ring

template <int EPI, bool PUSH>
__global__ void __launch_bounds__(ring::THREADS, 1)
    inv_ring_kernel(const __grid_constant__ CUtensorMap map_plane, const __grid_constant__ CUtensorMap map_nyq, const Tile* __restrict__ tiles,
                    EpilogueArgs ep, int bz, int zlo, int max_nz, int n_planes, int planes_per_cta, int interleave, Twiddles tw) {
    using namespace ring;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float2* slots = reinterpret_cast<float2*>(smem_raw);
    float2* twr = reinterpret_cast<float2*>(smem_raw + OFF_TW);
    float2* twu = twr + 64;
    float2* w128 = reinterpret_cast<float2*>(smem_raw + OFF_W128);
    PeerTarget* peers_all = reinterpret_cast<PeerTarget*>(smem_raw + OFF_PEER);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + OFF_BAR);
    __shared__ double s_sum[THREADS / 32];
    __shared__ float s_max[THREADS / 32];
    __shared__ int s_npeer[G];

    const int tid = threadIdx.x;
    const int g = tid / GT, lt = tid - g * GT;
    const int warp = lt >> 5, lane = lt & 31;
    row128_fill_tables(twr, twu, tw.t[6], tw.t[7], true, tid, THREADS);
    for (int i = tid; i < 128; i += THREADS) w128[i] = __ldg(&tw.t[7][i]);
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) tma::mbar_init(&full[s], 1);
        tma::fence_barrier_init();
    }
    if (tid < G) s_npeer[tid] = 0;
    __syncthreads();

    // plane k of this CTA = (tile, effective plane) pair q; slot k % S; group k % G.  interleave = 0: every CTA a contiguous
    // run of pairs; 1: pairs dealt round-robin, so that the CTAs in flight work on neighbouring planes of the same tile
    // (their epilogue rows share DRAM pages and the cache lines at tile seams)
    const int q0 = interleave ? (int)blockIdx.x : (int)blockIdx.x * planes_per_cta;
    const int qstep = interleave ? (int)gridDim.x : 1;  // interleave >= 2: additionally L2-prefetch the epilogue operands
    const int count = interleave ? (n_planes > q0 ? (n_planes - q0 + qstep - 1) / qstep : 0) : min(planes_per_cta, n_planes - q0);

    auto plane_of = [&](int k, int& ti, int& zr) {
        const int q = q0 + k * qstep;
        ti = q / max_nz;
        zr = q - ti * max_nz;
    };
    // executed by one thread.  Every (tile, plane) pair below the class maximum is fetched, also the few pairs beyond a
    // tile's own plane count (their slot is consumed without being transformed): the issue path reads no global memory.
    auto issue_load = [&](int k) {
        if (k >= count) return;
        int ti, zr;
        plane_of(k, ti, zr);
        const int row = ROWS_PER_PLANE * (ti * bz + zlo + zr);
        const int s = k % S;
        float2* slot = slots + (size_t)s * SLOT_ELEMS;
        tma::mbar_arrive_expect_tx(&full[s], PLANE_BYTES);
        tma::load_2d(slot, &map_plane, &full[s], 0, row);
        tma::load_2d(slot + BY * P, &map_nyq, &full[s], 0, row + BY);
    };
    if (tid == 0) {
        for (int k = 0; k < S; ++k) issue_load(k);
    }

    auto gsync = [&] { tma::group_sync(1 + g, GT); };
    const bool use_tik = ep.lambda > 0.0;
    const TikhonovConst tik = {ep.tik_hi, ep.tik_lo, ep.lambda};
    double lsum = 0.0;
    float lmax = 0.f;
    PeerTarget* s_peer = peers_all + g * kMaxPeers;

    // column phase mapping: four columns per warp, eight threads per column
    const int cq = lane & 3, t = lane >> 2;
    const int sw = t & 3;
    int last_ti = -1;
    Tile tile;

#pragma unroll 1
    for (int k = g; k < count; k += G) {
        int ti, zr;
        plane_of(k, ti, zr);
        if (ti != last_ti) {
            tile = tiles[ti];
            last_ti = ti;
            if constexpr (PUSH) {
                // the group's previous plane is behind the barrier that ended its row phase
                if (lt < 32) {
                    bool hit = false;
                    PeerTarget pt;
                    if (lt < ep.n_peers) {
                        pt = ep.peers[lt];
                        hit = true;
#pragma unroll
                        for (int d = 0; d < 3; ++d) hit = hit && tile.e[d] < pt.hi[d] && tile.e[d] + tile.n[d] > pt.lo[d];
                    }
                    const unsigned mk = __ballot_sync(0xffffffffu, hit);
                    if (hit) s_peer[__popc(mk & ((1u << lt) - 1u))] = pt;
                    if (lt == 0) s_npeer[g] = __popc(mk);
                }
                // read by the row phase behind the barrier that ends the column phase
            }
        }
        const bool valid = zr < tile.n[2];  // otherwise the pair lies beyond the tile's own planes: consume the slot only
        const int s = k % S;
        float2* __restrict__ plane = slots + (size_t)s * SLOT_ELEMS;
        const float2* __restrict__ nyq = plane + BY * P;
        const int zv = tile.e[2] + zr;  // plane inside the volume

        if (valid && EPI != EPI_STORE && interleave >= 2) {
            // the epilogue's operand rows of this plane towards L2 while the plane's spectrum is still in flight / being
            // transformed: one 128-byte line per thread and step, 4 lines cover a 98-float row
            const int n0p = tile.n[0], n1p = tile.n[1];
            for (int i = lt; i < 4 * n1p; i += GT) {
                const int yl = i >> 2, seg = i & 3;
                if (32 * seg < n0p) {
                    const size_t ob = ((size_t)(zv + ep.out_off[2]) * ep.out_dim[1] + (tile.e[1] + yl + ep.out_off[1])) * (size_t)ep.out_dim[0] +
                                      tile.e[0] + ep.out_off[0] + 32 * seg;
                    const size_t ab = ((size_t)zv * ep.aux_dim[1] + (tile.e[1] + yl)) * (size_t)ep.aux_dim[0] + tile.e[0] + 32 * seg;
                    if constexpr (EPI == EPI_QUOTIENT) asm volatile("prefetch.global.L2 [%0];" ::"l"(ep.img + ab));
                    if constexpr (EPI == EPI_UPDATE) {
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(ep.out + ob));
                        if (ep.weight) asm volatile("prefetch.global.L2 [%0];" ::"l"(ep.weight + ab));
                    }
                }
            }
        }
        tma::mbar_wait(&full[s], (unsigned)(k / S) & 1u);
        if (valid) {
        // ---- phase 1: inverse columns, in place, warp-local; column 0 re-packs DC + i*Nyquist
#pragma unroll 1
        for (int cr = 0; cr < H / 32; ++cr) {
            const int col = cr * 32 + warp * 4 + cq;
            float2* __restrict__ cp = plane + col;
            float2 v[16];
#pragma unroll
            for (int m = 0; m < 16; ++m) v[m] = cp[(t + 8 * m) * P];
            if (col == 0) {
#pragma unroll
                for (int m = 0; m < 16; ++m) {
                    const float2 b = nyq[t + 8 * m];
                    v[m] = make_float2(v[m].x - b.y, v[m].y + b.x);
                }
            }
            __syncwarp();  // every thread of the column holds its points before the in-place scatter
            // pass 1: radix 16 (no twiddles), X1[r] of sub-sequence t -> position 16 t + r, stored in row
            // 16 t + (r & ~3) + ((r & 3) ^ (t & 3)): the four columns of a thread quad and the four t of a half-warp
            // cover the 32 banks exactly once (row pitch 68 = 8 banks mod 32)
            dft16<true>(v);
            {
                float2* const pw0 = cp + (16 * t + (0 ^ sw)) * P;
                float2* const pw1 = cp + (16 * t + (1 ^ sw)) * P;
                float2* const pw2 = cp + (16 * t + (2 ^ sw)) * P;
                float2* const pw3 = cp + (16 * t + (3 ^ sw)) * P;
#pragma unroll
                for (int r = 0; r < 16; ++r) {
                    float2* pw = (r & 3) == 0 ? pw0 : ((r & 3) == 1 ? pw1 : ((r & 3) == 2 ? pw2 : pw3));
                    pw[(r & ~3) * P] = v[r];
                }
            }
            __syncwarp();
            // gather position t + 8 m = row 8 m + (t ^ ((m >> 1) & 3))
            {
                const float2* const pg0 = cp + (t ^ 0) * P;
                const float2* const pg1 = cp + (t ^ 1) * P;
                const float2* const pg2 = cp + (t ^ 2) * P;
                const float2* const pg3 = cp + (t ^ 3) * P;
#pragma unroll
                for (int m = 0; m < 16; ++m) {
                    const int j = (m >> 1) & 3;
                    const float2* pg = j == 0 ? pg0 : (j == 1 ? pg1 : (j == 2 ? pg2 : pg3));
                    v[m] = pg[(8 * m) * P];
                }
            }
            __syncwarp();  // all gathers done before the natural-order results overwrite the swizzled rows
            // pass 2: radix 8 with the inter-pass twiddles conj(w128^(k r)), k = t + 8 b, in registers
            auto noput = [](int, float2) {};
            fft_pass<128, 16, 8, 16, true, true>(v, t, w128, noput);
#pragma unroll
            for (int m = 0; m < 16; ++m) cp[(t + 8 * m) * P] = v[m];
        }
        gsync();  // all columns of the plane are transformed

        // ---- phase 2: inverse rows of the effective region + epilogue
        inv_rows128<EPI, PUSH, GT>(plane, twr, twu, tile, zv, ep, s_peer, PUSH ? s_npeer[g] : 0, lt, use_tik, tik, lsum, lmax);
        }
        if constexpr (PUSH) {
            if (valid && s_npeer[g]) {  // the same for every thread of the group: interior tiles pay nothing
                gsync();  // the plane's rows hold the results
                push_plane128<GT>(plane, tile, zv, s_peer, s_npeer[g], lt);
            }
        }
        tma::fence_proxy_async();  // this thread's accesses to the slot are ordered before the bulk copies that refill it
        gsync();  // the slot is consumed: refill it with the plane three ahead
        if (lt == 0) issue_load(k + S);
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
                double sm = 0.0;
                float mx = 0.f;
                for (int i = 0; i < THREADS / 32; ++i) { sm += s_sum[i]; mx = fmaxf(mx, s_max[i]); }
                atomicAdd(ep.sum_change, sm);
                atomicMax(ep.max_change_bits, __float_as_uint(mx));
            }
        }
    }
}

static inline Twiddles to_dev(const TwiddleTables& t) {
    Twiddles d;
    for (int i = 0; i < 11; ++i) d.t[i] = t.t[i];
    return d;
}

bool inv_ring_supported(int bx, int by) { return bx == 128 && by == 128; }

cudaError_t configure_inv_ring_kernels() {
    cudaError_t e = cudaSuccess;
#define CFG(EPI, PUSH) \
    if (e == cudaSuccess) e = cudaFuncSetAttribute(inv_ring_kernel<EPI, PUSH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring::SMEM);
    CFG(EPI_STORE, false) CFG(EPI_QUOTIENT, false) CFG(EPI_UPDATE, false) CFG(EPI_QUOTIENT, true) CFG(EPI_UPDATE, true)
#undef CFG
    return e;
}

cudaError_t launch_inv_ring(int bz, int epilogue, const float2* scratch, const Tile* d_tiles, int n_tiles, int zlo, int max_nz,
                            const EpilogueArgs& ep, int num_sms, int interleave, const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0 || max_nz <= 0) return cudaSuccess;
    if (ep.n_peers > kMaxPeers || zlo + max_nz > bz) return cudaErrorInvalidValue;
    const long long total = (long long)n_tiles * max_nz;
    const unsigned long long rows = (unsigned long long)n_tiles * bz * ring::ROWS_PER_PLANE;
    if (total > 0x7fffffffLL || rows > 0x7fffffffULL) return cudaErrorNotSupported;
    const int n_planes = (int)total;
    // the batch's scratch as rows of 64 complex values (128 floats): a plane is 128 rows + 2 rows holding its Nyquist column
    CUtensorMap mp, mn;
    const unsigned long long dims[2] = {2ull * ring::H, rows};
    const unsigned long long strides[1] = {2ull * ring::H * sizeof(float)};
    const unsigned box_plane[2] = {2u * ring::P, (unsigned)ring::BY};  // 68 complex wide: the 4 columns beyond the row are zero-filled
    const unsigned box_nyq[2] = {2u * ring::H, 2u};
    if (!encode_tensor_map_f32(&mp, scratch, 2, dims, strides, box_plane) || !encode_tensor_map_f32(&mn, scratch, 2, dims, strides, box_nyq))
        return cudaErrorNotSupported;
    // every CTA a contiguous run of planes, at least a handful so that the ring pays off
    const int grid = std::max(1, std::min(num_sms, (n_planes + 5) / 6));
    const int per = (n_planes + grid - 1) / grid;
    const bool push = ep.peers != nullptr && ep.n_peers > 0;
#define L(EPI, PUSH) inv_ring_kernel<EPI, PUSH><<<grid, ring::THREADS, ring::SMEM, st>>>(mp, mn, d_tiles, ep, bz, zlo, max_nz, n_planes, per, interleave, to_dev(tw))
    if (epilogue == EPI_STORE) L(EPI_STORE, false);
    else if (epilogue == EPI_QUOTIENT && !push) L(EPI_QUOTIENT, false);
    else if (epilogue == EPI_QUOTIENT) L(EPI_QUOTIENT, true);
    else if (!push) L(EPI_UPDATE, false);
    else L(EPI_UPDATE, true);
#undef L
    return cudaGetLastError();
}

}  // namespace mvd
