// z pass of the tiled FFT convolution with TMA data movement (sm_100a).
//
// Same arithmetic and thread mapping as z_pass_kernel<BZ, 1> (conv_kernels.cu): forward z FFT, multiply with
// the resident kernel spectrum, inverse z FFT on work items of PW = 16 plane positions x BZ planes, lanes across
// positions.  What changes is how an item moves:
//
//   * one elected thread fetches the whole item (a PW x BZ box of the batch's [tile][z][position] scratch, 128 B x BZ
//     rows that lie plane_elems * 8 bytes apart) with cp.async.bulk.tensor into a shared-memory slot and the
//     transform runs IN the slot (the box layout [z][PW] is the exchange layout of the line FFT);
//   * results go back into the slot and leave with one bulk tensor store; rows at or beyond the last plane
//     inv_xy consumes are clipped by the store map's extent;
//   * a CTA holds S = 3 slots and G = 2 independent thread groups (named barriers): while the groups transform
//     two items, the third slot is in flight -- the load of item k + 3 is issued the moment item k's store has
//     read its slot.  No thread computes a global address or holds a load in a register.
//
// Replaces the z part of J/fftMethods/FFTConvolution.java:470-551 (C2C along z, multiply :607-614, inverse).
#include "conv_kernels.h"
#include "fft_device.cuh"
#include "tma.cuh"
#include "z_items.cuh"

#include <algorithm>

namespace mvd {

template <int BZ> struct ZT {
    static constexpr int EPT = BZ >= 128 ? 16 : 8;
    static constexpr int TZ = BZ / EPT;
    static constexpr int PW = 16;
    static constexpr int GT = PW * TZ;  // threads per group
    static constexpr int G = 2;         // groups per CTA
    static constexpr int S = 3;         // slots per CTA
    static constexpr int THREADS = G * GT;
    static constexpr int BOXZ = BZ < 256 ? BZ : 256;  // a box dimension holds at most 256 elements
    static constexpr int NBOX = BZ / BOXZ;
    static constexpr unsigned SLOT_BYTES = (unsigned)PW * BZ * sizeof(float2);
    // slots | twiddles | mbarriers
    static constexpr size_t SMEM = (size_t)S * SLOT_BYTES + (size_t)BZ * sizeof(float2) + 64;
    static constexpr int MIN_CTAS_ = (int)(232448 / (SMEM + 1024));
    static constexpr int MIN_CTAS = MIN_CTAS_ < 1 ? 1 : (MIN_CTAS_ * THREADS > 1024 ? 1024 / THREADS : MIN_CTAS_);
};

template <int BZ>
__global__ void __launch_bounds__(ZT<BZ>::THREADS, ZT<BZ>::MIN_CTAS)
    z_tma_kernel(const __grid_constant__ CUtensorMap map_ld, const __grid_constant__ CUtensorMap map_st, const float2* __restrict__ kspec,
                 int plane_elems, int n_tiles, int n_items, Twiddles tw) {
    using C = ZT<BZ>;
    constexpr int EPT = C::EPT, TZ = C::TZ, PW = C::PW, GT = C::GT, G = C::G, S = C::S;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float2* slots = reinterpret_cast<float2*>(smem_raw);
    float2* twz = reinterpret_cast<float2*>(smem_raw + (size_t)S * C::SLOT_BYTES);
    uint64_t* full = reinterpret_cast<uint64_t*>(twz + BZ);

    const int tid = threadIdx.x;
    const int g = tid / GT, lt = tid - g * GT;
    const int p = lt % PW, t = lt / PW;
    for (int i = tid; i < BZ; i += C::THREADS) twz[i] = __ldg(&tw.t[ilog2<BZ>()][i]);
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) tma::mbar_init(&full[s], 1);
        tma::fence_barrier_init();
    }
    __syncthreads();

    // item k of this CTA = work item blockIdx.x + k * gridDim.x, lives in slot k % S, transformed by group k % G
    auto issue_load = [&](int k) {
        const int item = blockIdx.x + k * gridDim.x;
        if (item >= n_items) return;
        int pblk, ti;
        decode_item(item, n_tiles, pblk, ti);
        const int s = k % S;
        tma::mbar_arrive_expect_tx(&full[s], C::SLOT_BYTES);
#pragma unroll
        for (int b = 0; b < C::NBOX; ++b)
            tma::load_3d(slots + (size_t)s * (PW * BZ) + (size_t)b * C::BOXZ * PW, &map_ld, &full[s], 2 * PW * pblk, b * C::BOXZ, ti);
    };
    if (tid == 0) {
#pragma unroll
        for (int k = 0; k < S; ++k) issue_load(k);
    }

    auto gsync = [&] { tma::group_sync(1 + g, GT); };
#pragma unroll 1
    for (int k = g;; k += G) {
        const int item = blockIdx.x + k * gridDim.x;
        if (item >= n_items) break;
        const int s = k % S;
        int pblk, ti;
        decode_item(item, n_tiles, pblk, ti);
        const int pos = pblk * PW + p;
        const bool valid = pos < plane_elems;
        const float2* __restrict__ kb = kspec + (valid ? pos : 0);
        float2* __restrict__ xp = slots + (size_t)s * (PW * BZ) + p;

        tma::mbar_wait(&full[s], (unsigned)(k / S) & 1u);
        float2 v[EPT];
#pragma unroll
        for (int m = 0; m < EPT; ++m) v[m] = xp[(t + m * TZ) * PW];
        gsync();  // everyone holds its points before the in-place exchange overwrites the slot

        auto put = [&](int idx, float2 val) { xp[idx * PW] = val; };
        auto get = [&](int idx) { return xp[idx * PW]; };
        line_fft<BZ, EPT, false>(v, t, twz, put, get, gsync);

        // spectrum values in batches (all EPT at once would spill: the compiler hoists every load above the first multiplication)
        constexpr int KB = BZ >= 512 ? 4 : 16;
#pragma unroll
        for (int h = 0; h < EPT; h += KB) {
            float2 kk[KB];
            const unsigned pl = (unsigned)plane_elems;
#pragma unroll
            for (int i = 0; i < KB; ++i) kk[i] = __ldg(kb + (unsigned)(t + (h + i) * TZ) * pl);
#pragma unroll
            for (int i = 0; i < KB; ++i) v[h + i] = cmul(v[h + i], kk[i]);
            asm volatile("" ::: "memory");
        }
        gsync();  // the forward transform's last gathers are complete before the inverse scatters
        line_fft<BZ, EPT, true>(v, t, twz, put, get, gsync);
        // a thread's results go where its own last gather read: no barrier needed before these stores
#pragma unroll
        for (int m = 0; m < EPT; ++m) xp[(t + m * TZ) * PW] = v[m];
        tma::fence_proxy_async();
        gsync();
        if (lt == 0) {
#pragma unroll
            for (int b = 0; b < C::NBOX; ++b)
                tma::store_3d(&map_st, slots + (size_t)s * (PW * BZ) + (size_t)b * C::BOXZ * PW, 2 * PW * pblk, b * C::BOXZ, ti);
            tma::store_commit();
            tma::store_wait_read();  // the slot has been read: refill it with the item three ahead
            issue_load(k + S);
        }
    }
    if (lt == 0) tma::store_wait_all();
}

// ---------------------------------------------------------------------------------------------
static bool encode_tensor_map(CUtensorMap* out, const void* base, int rank, const unsigned long long* dims, const unsigned long long* strides_bytes,
                              const unsigned* box, bool swizzle128) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p)
            return false;
        fn = reinterpret_cast<EncodeFn>(p);
    }
    cuuint64_t gd[5], gs[5];
    cuuint32_t bd[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gd[i] = dims[i];
        bd[i] = box[i];
        es[i] = 1;
        if (i + 1 < rank) gs[i] = strides_bytes[i];
    }
    return fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bd, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
bool encode_tensor_map_f32(CUtensorMap* out, const void* base, int rank, const unsigned long long* dims, const unsigned long long* strides_bytes,
                           const unsigned* box) {
    return encode_tensor_map(out, base, rank, dims, strides_bytes, box, false);
}
// 128-byte swizzle: the box's inner extent must be 128 bytes and the shared-memory destination 1024-byte aligned
bool encode_tensor_map_f32_swizzle128(CUtensorMap* out, const void* base, int rank, const unsigned long long* dims,
                                      const unsigned long long* strides_bytes, const unsigned* box) {
    return encode_tensor_map(out, base, rank, dims, strides_bytes, box, true);
}

bool z_tma_supported(int bz) { return bz == 128 || bz == 256 || bz == 512; }

static inline Twiddles to_dev(const TwiddleTables& t) {
    Twiddles d;
    for (int i = 0; i < 11; ++i) d.t[i] = t.t[i];
    return d;
}

cudaError_t configure_z_tma_kernels() {
    cudaError_t e = cudaSuccess;
#define CFG(BZ) \
    if (e == cudaSuccess) e = cudaFuncSetAttribute(z_tma_kernel<BZ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZT<BZ>::SMEM);
    CFG(128) CFG(256) CFG(512)
#undef CFG
    return e;
}

// zlo = first plane of a tile that inv_xy consumes, max_nz = the most planes any tile of the batch hands on
cudaError_t launch_z_pass_tma(int bx, int by, int bz, float2* scratch, const float2* kspec, int n_tiles, int zlo, int max_nz, int num_sms,
                              const TwiddleTables& tw, cudaStream_t st) {
    if (n_tiles <= 0) return cudaSuccess;
    if (!z_tma_supported(bz)) return cudaErrorInvalidValue;
    const unsigned long long pe = plane_elems(bx, by);
    const int zext = std::min(bz, zlo + max_nz);
    CUtensorMap ld, stm;
    const unsigned long long strides[2] = {pe * 8ull, pe * 8ull * (unsigned long long)bz};
    const unsigned long long dims_ld[3] = {2ull * pe, (unsigned long long)bz, (unsigned long long)n_tiles};
    const unsigned long long dims_st[3] = {2ull * pe, (unsigned long long)zext, (unsigned long long)n_tiles};
    const unsigned box[3] = {32u, (unsigned)std::min(bz, 256), 1u};
    if (!encode_tensor_map_f32(&ld, scratch, 3, dims_ld, strides, box) || !encode_tensor_map_f32(&stm, scratch, 3, dims_st, strides, box))
        return cudaErrorInvalidValue;
    const int n_pblk = (int)((pe + 15) / 16);
    const int n_items = n_tiles * ((n_pblk + MVD_Z_PGROUP - 1) / MVD_Z_PGROUP * MVD_Z_PGROUP);
#define F(BZ)                                                                                                                   \
    {                                                                                                                           \
        const int cap = num_sms * ZT<BZ>::MIN_CTAS;                                                                             \
        const int grid = std::max(1, std::min((n_items + ZT<BZ>::G - 1) / ZT<BZ>::G, cap));                                     \
        z_tma_kernel<BZ><<<grid, ZT<BZ>::THREADS, ZT<BZ>::SMEM, st>>>(ld, stm, kspec, (int)pe, n_tiles, n_items, to_dev(tw));    \
    }
    switch (bz) {
        case 128: F(128); break;
        case 256: F(256); break;
        default: F(512); break;
    }
#undef F
    return cudaGetLastError();
}

}  // namespace mvd
