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
// 5 K instructions per voxel instead of 250.  Both kernels march SEP_U rows / planes per step so that
// several loads are in flight and the register window shifts once per SEP_U outputs.  The same kernels serve the reference's second native
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
#ifndef MVD_SEP_CHY
#define MVD_SEP_CHY 256
#endif
constexpr int SEP_CHY = MVD_SEP_CHY;  // rows per CTA in the xy kernel: the x pass also runs on the KMAX - 1 halo rows of a chunk
#ifndef MVD_SEP_CHZ
#define MVD_SEP_CHZ 256
#endif
constexpr int SEP_CHZ = MVD_SEP_CHZ;  // planes per CTA in the z kernel (+ KMAX - 1 halo planes read per chunk)

constexpr int SEP_R = 16;    // rows per marching step of the xy kernel
constexpr int SEP_WP = 164;  // pitch of the input strip: >= 128 + 31 - 1, multiple of 4, = 4 (mod 32) -> conflict-free LDS.128 across 8 rows
constexpr int SEP_XP = 132;  // pitch of the x-convolved strip, same rule
constexpr int SEP_U = 4;     // planes per marching step of the z kernel

template <int KMAX>
__global__ void __launch_bounds__(SEP_NT) sep_xy_kernel(Volume src, SepTaps taps, float* __restrict__ tmp, int bx, int by) {
    // tmp: [src.dim[2]][by][bx] -- every resident plane of the source (brick and ghost planes)
    constexpr int H = KMAX / 2, R = SEP_R, W = SEP_NT + KMAX - 1;
    constexpr int NIN = (16 + KMAX - 1 + 3) / 4;  // float4 loads feeding 16 adjacent outputs
    constexpr int HW = KMAX - 1;                  // halo columns of the strip (right of the CTA's 128)
    constexpr int NH = (HW * R + SEP_NT - 1) / SEP_NT;  // halo elements per thread and step
    static_assert(W <= SEP_WP && 4 * NIN + 16 * 7 <= SEP_WP, "input strip too narrow");
    __shared__ __align__(16) float in[R][SEP_WP];
    __shared__ __align__(16) float xs[R][SEP_XP];
    __shared__ int xmap[W];
    __shared__ int ymap[2][R];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * SEP_NT, y0 = blockIdx.y * SEP_CHY, za = blockIdx.z;
    const float fillv = src.ext == EXT_CONST ? src.ext_value : 0.f;
    for (int i = tid; i < W; i += SEP_NT) xmap[i] = map_coord(x0 - H + i, src.dim[0], src.off[0], src.gmin[0], src.gdim[0], src.ext);
    for (int i = tid; i < R * SEP_WP; i += SEP_NT) (&in[0][0])[i] = 0.f;  // the pad columns are read (times zero taps): keep them finite
    float sx = 0.f;  // x convolution of a constant row = fill * sum of the x taps
#pragma unroll
    for (int i = 0; i < KMAX; ++i) sx += taps.x[i];
    const float fill_row = taps.per_pass_fill ? fillv : fillv * sx;
    const float* __restrict__ zbase = src.data + (size_t)za * src.dim[1] * (size_t)src.dim[0];
    // win[j] = x-convolved row (r0 - (KMAX - 1) + j) of this thread's column; output row y = r0 + u - H reads win[u .. u + KMAX - 1]
    float win[KMAX - 1 + R];
#pragma unroll
    for (int i = 0; i < KMAX - 1 + R; ++i) win[i] = 0.f;
    const int yend = min(y0 + SEP_CHY, by);
    // x pass mapping: 8 consecutive threads take 8 different rows (conflict-free float4 reads), a thread makes 16 adjacent outputs
    const int xr = (tid & 7) + 8 * ((tid >> 3) & 1), xseg = (tid >> 4) * 16;
    // The strip of the NEXT step is requested before the current step is convolved and waits in registers
    // (pre: this thread's column of the 16 rows, coalesced; preh: its share of the halo columns), so the load
    // latency hides behind the 2 x 16 x KMAX multiply-adds of a step.
    float pre[R], preh[NH];
    if (tid < R) ymap[0][tid] = map_coord(y0 - H + tid, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
    __syncthreads();
    const int xm = xmap[tid];
    auto sample = [&](int yl, int xi) -> float {
        return yl >= 0 ? (xi >= 0 ? __ldg(zbase + (size_t)yl * src.dim[0] + xi) : (xi == -1 ? fillv : 0.f)) : 0.f;
    };
    auto request = [&](const int* ym) {
#pragma unroll
        for (int r = 0; r < R; ++r) pre[r] = sample(ym[r], xm);
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const int e = tid + SEP_NT * j;
            preh[j] = 0.f;
            if (e < HW * R) {
                const int r = e / HW;
                preh[j] = sample(ym[r], xmap[SEP_NT + e - r * HW]);
            }
        }
    };
    request(ymap[0]);
    int cur = 0;
#pragma unroll 1
    for (int r0 = y0 - H; r0 < yend + H; r0 += R, cur ^= 1) {
        __syncthreads();  // the previous step's column reads of xs are done, ymap[cur ^ 1] is free
#pragma unroll
        for (int r = 0; r < R; ++r) in[r][tid] = pre[r];
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const int e = tid + SEP_NT * j;
            if (e < HW * R) {
                const int r = e / HW;
                in[r][SEP_NT + e - r * HW] = preh[j];
            }
        }
        if (tid < R) ymap[cur ^ 1][tid] = map_coord(r0 + R + tid, src.dim[1], src.off[1], src.gmin[1], src.gdim[1], src.ext);
        __syncthreads();
        if (r0 + R < yend + H) request(ymap[cur ^ 1]);
        const int* __restrict__ ym = ymap[cur];
        if (ym[xr] >= 0) {
            float v[4 * NIN];
#pragma unroll
            for (int j = 0; j < NIN; ++j) {
                const float4 q = *reinterpret_cast<const float4*>(&in[xr][xseg + 4 * j]);
                v[4 * j] = q.x; v[4 * j + 1] = q.y; v[4 * j + 2] = q.z; v[4 * j + 3] = q.w;
            }
#pragma unroll
            for (int o4 = 0; o4 < 4; ++o4) {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int i = 0; i < KMAX; ++i) {
#pragma unroll
                    for (int o = 0; o < 4; ++o) acc[o] = fmaf(taps.x[i], v[4 * o4 + o + i], acc[o]);
                }
                *reinterpret_cast<float4*>(&xs[xr][xseg + 4 * o4]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < KMAX - 1; ++i) win[i] = win[i + R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int yl = ym[r];
            win[KMAX - 1 + r] = yl >= 0 ? xs[r][tid] : (yl == -1 ? fill_row : 0.f);
        }
        float* __restrict__ op = tmp + ((size_t)za * by + (r0 - H)) * bx + x0 + tid;
#pragma unroll
        for (int u = 0; u < R; ++u) {
            const int y = r0 + u - H;
            if (y >= y0 && y < yend && x0 + tid < bx) {
                float acc = 0.f;
#pragma unroll
                for (int i = 0; i < KMAX; ++i) acc = fmaf(taps.y[i], win[u + i], acc);
                op[(size_t)u * bx] = acc;
            }
        }
    }
}

// PUSH: fused ghost exchange as in the FFT path's epilogue -- every output voxel a neighbouring brick needs is
// also stored into that neighbour's ghost layers (PeerTarget, conv_types.h).  A thread owns one (x, y) column,
// so the neighbours whose x / y range contains it are found once; per voxel only the z range is tested.
template <int KMAX, int EPI, bool PUSH>
__global__ void __launch_bounds__(SEP_NT) sep_z_kernel(const float* __restrict__ tmp, Volume src, SepTaps taps, EpilogueArgs ep, int bx,
                                                       int by, int bz) {
    constexpr int H = KMAX / 2, U = SEP_U;
    __shared__ double s_sum[SEP_NT / 32];
    __shared__ float s_max[SEP_NT / 32];
    __shared__ PeerTarget s_peer[PUSH ? kMaxPeers : 1];
    const int tid = threadIdx.x;
    const int x = blockIdx.x * SEP_NT + tid, y = blockIdx.y, z0 = blockIdx.z * SEP_CHZ;
    const bool live = x < bx;
    unsigned pmask = 0;
    if constexpr (PUSH) {
        for (int i = tid; i < ep.n_peers; i += SEP_NT) s_peer[i] = ep.peers[i];
        __syncthreads();
        if (live)
            for (int i = 0; i < ep.n_peers; ++i) {
                const PeerTarget& t = s_peer[i];
                if (x >= t.lo[0] && x < t.hi[0] && y >= t.lo[1] && y < t.hi[1]) pmask |= 1u << i;
            }
    }
    auto push = [&](int z, float val) {
        if constexpr (PUSH) {
            for (unsigned m = pmask; m; m &= m - 1) {
                const PeerTarget& t = s_peer[__ffs(m) - 1];
                if (z >= t.lo[2] && z < t.hi[2])
                    t.base[((size_t)(z + t.shift[2]) * t.dim[1] + (y + t.shift[1])) * (size_t)t.dim[0] + (x + t.shift[0])] = val;
            }
        }
    };
    const float fillv = src.ext == EXT_CONST ? src.ext_value : 0.f;
    float sxy = 0.f, sy = 0.f;  // a plane of constant fill convolved in x and y
#pragma unroll
    for (int i = 0; i < KMAX; ++i) { sxy += taps.x[i]; sy += taps.y[i]; }
    const float fill_t = taps.per_pass_fill ? fillv : fillv * sxy * sy;
    const bool use_tik = ep.lambda > 0.0;
    const TikhonovConst tik = {ep.tik_hi, ep.tik_lo, ep.lambda};
    double lsum = 0.0;
    float lmax = 0.f;
    float win[KMAX - 1 + U];  // win[j] = plane (r0 - (KMAX - 1) + j) of tmp at this thread's (x, y)
#pragma unroll
    for (int i = 0; i < KMAX - 1 + U; ++i) win[i] = 0.f;
    const int zend = min(z0 + SEP_CHZ, bz);
    const size_t plane = (size_t)by * bx;
    const float* __restrict__ tcol = tmp + (size_t)y * bx + (live ? x : 0);
    const size_t oplane = (size_t)ep.out_dim[1] * ep.out_dim[0], aplane = (size_t)ep.aux_dim[1] * ep.aux_dim[0];
    // pointers of output plane z = r0 - H, advanced by U planes per step
    float* optr = ep.out + ((size_t)(z0 - 2 * H + ep.out_off[2]) * ep.out_dim[1] + (y + ep.out_off[1])) * (size_t)ep.out_dim[0] + ep.out_off[0] + (live ? x : 0);
    const float* iptr = ep.img + ((size_t)(z0 - 2 * H) * ep.aux_dim[1] + y) * (size_t)ep.aux_dim[0] + (live ? x : 0);
    const float* wptr = ep.weight + ((size_t)(z0 - 2 * H) * ep.aux_dim[1] + y) * (size_t)ep.aux_dim[0] + (live ? x : 0);
    // One step ahead: the U planes of tmp and the epilogue operands of the NEXT step are requested before the
    // current step's multiply-adds and wait in registers (nxt, an, bn).
    float nxt[U], an[U], bn[U];
    auto request = [&](int r0n, const float* op, const float* ip, const float* wp) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int zl = map_coord(r0n + u, src.dim[2], src.off[2], src.gmin[2], src.gdim[2], src.ext);
            nxt[u] = zl >= 0 ? __ldg(tcol + (size_t)zl * plane) : (zl == -1 ? fill_t : 0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int z = r0n + u - H;
            an[u] = 0.f;
            bn[u] = ep.weight_const;
            if (z >= z0 && z < zend && live) {
                if constexpr (EPI == EPI_QUOTIENT) an[u] = __ldg(ip + u * aplane);
                if constexpr (EPI == EPI_UPDATE) {
                    an[u] = op[u * oplane];
                    if (ep.weight) bn[u] = __ldg(wp + u * aplane);
                }
            }
        }
    };
    request(z0 - H, optr, iptr, wptr);
#pragma unroll 1
    for (int r0 = z0 - H; r0 < zend + H; r0 += U, optr += U * oplane, iptr += U * aplane, wptr += U * aplane) {
#pragma unroll
        for (int i = 0; i < KMAX - 1; ++i) win[i] = win[i + U];
        float a[U], b[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            win[KMAX - 1 + u] = nxt[u];
            a[u] = an[u];
            b[u] = bn[u];
        }
        if (r0 + U < zend + H) request(r0 + U, optr + U * oplane, iptr + U * aplane, wptr + U * aplane);
        float acc[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int z = r0 + u - H;
            ok[u] = z >= z0 && z < zend && live;
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < KMAX; ++i) s = fmaf(taps.z[i], win[u + i], s);
            acc[u] = s + ep.add_const;
        }
        if constexpr (EPI == EPI_UPDATE) {
            // the step's four voxels together (next_psi_quad: no branch per voxel, device_common.cuh); voxels outside the
            // chunk carry a = 0 and are neither stored nor counted; their change is summed in float first (four terms)
            static_assert(U == 4, "the update epilogue takes the planes of a step as one quad");
            float nv[4];
            next_psi_quad(a, acc, b, use_tik, tik, ep.min_value, nv);
            float rs = 0.f;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!ok[u]) continue;
                optr[u * oplane] = nv[u];
                push(r0 + u - H, nv[u]);
                const float ch = fabsf(__fsub_rn(nv[u], a[u]));
                rs += ch;
                lmax = fmaxf(lmax, ch);
            }
            lsum += (double)rs;
        } else {
            float ov[U];
            if constexpr (EPI == EPI_QUOTIENT) {
                static_assert(U == 4, "the quotient epilogue takes the planes of a step as one quad");
                quotient_quad(a, acc, ov);  // voxels outside the chunk carry a = 0: no division, not stored
            } else {
#pragma unroll
                for (int u = 0; u < U; ++u) ov[u] = acc[u];
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!ok[u]) continue;
                optr[u * oplane] = ov[u];
                push(r0 + u - H, ov[u]);
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
    const bool push = ep.peers != nullptr && ep.n_peers > 0;
    if (push && (ep.n_peers > kMaxPeers || epilogue == EPI_STORE)) return cudaErrorInvalidValue;
#define SEP_LAUNCH(KM)                                                                                                  \
    {                                                                                                                   \
        sep_xy_kernel<KM><<<g1, SEP_NT, 0, st>>>(src, taps, tmp, bx, by);                                               \
        if (epilogue == EPI_STORE) sep_z_kernel<KM, EPI_STORE, false><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz); \
        else if (epilogue == EPI_QUOTIENT && !push) sep_z_kernel<KM, EPI_QUOTIENT, false><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz); \
        else if (epilogue == EPI_QUOTIENT) sep_z_kernel<KM, EPI_QUOTIENT, true><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz); \
        else if (!push) sep_z_kernel<KM, EPI_UPDATE, false><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz);      \
        else sep_z_kernel<KM, EPI_UPDATE, true><<<g2, SEP_NT, 0, st>>>(tmp, src, taps, ep, bx, by, bz);                  \
    }
    if (kmax == 7) SEP_LAUNCH(7)
    else if (kmax == 15) SEP_LAUNCH(15)
    else if (kmax == 31) SEP_LAUNCH(31)
    else return cudaErrorInvalidValue;
#undef SEP_LAUNCH
    return cudaGetLastError();
}

}  // namespace mvd
