#include "volume_kernels.h"

namespace mvd {

// grid-stride kernels: 8 CTAs per SM of the current device (148 on B200)
static int sm_count() {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
    return n;
}

__global__ void fill_kernel(float* __restrict__ p, size_t n, float v) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

cudaError_t launch_fill(float* p, size_t n, float v, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const int blocks = (int)((n + 255) / 256 < (size_t)(sm_count() * 8) ? (n + 255) / 256 : (size_t)(sm_count() * 8));
    fill_kernel<<<blocks, 256, 0, st>>>(p, n, v);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) first_iteration_kernel(ViewPtrs views, size_t n, double* __restrict__ d_sum,
                                                              unsigned long long* __restrict__ d_count) {
    __shared__ double s_sum[8];
    __shared__ unsigned long long s_cnt[8];
    double lsum = 0.0;
    unsigned long long lcnt = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        double sum = 0.0;
        int count = 0;
        for (int v = 0; v < views.n; ++v) {
            const double x = (double)__ldg(views.img[v] + i);
            if (x > 0) { sum += x; ++count; }
        }
        if (count > 0) { lsum += sum / (double)count; ++lcnt; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
        lcnt += __shfl_xor_sync(0xffffffffu, lcnt, o);
    }
    if ((threadIdx.x & 31) == 0) { s_sum[threadIdx.x >> 5] = lsum; s_cnt[threadIdx.x >> 5] = lcnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        unsigned long long c = 0;
        for (int i = 0; i < 8; ++i) { s += s_sum[i]; c += s_cnt[i]; }
        atomicAdd(d_sum, s);
        atomicAdd(d_count, c);
    }
}

cudaError_t launch_first_iteration_sums(const ViewPtrs& views, size_t n, double* d_sum, unsigned long long* d_count, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const int blocks = (int)((n + 255) / 256 < (size_t)(sm_count() * 8) ? (n + 255) / 256 : (size_t)(sm_count() * 8));
    first_iteration_kernel<<<blocks, 256, 0, st>>>(views, n, d_sum, d_count);
    return cudaGetLastError();
}

struct Int3 { int v[3]; };

__global__ void __launch_bounds__(256) mask_kernel(ViewPtrs views, Int3 brick, float* __restrict__ psi, Int3 pdim, Int3 poff) {
    const size_t n = (size_t)brick.v[0] * brick.v[1] * brick.v[2];
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        bool any = false;
        for (int v = 0; v < views.n; ++v) any |= __ldg(views.img[v] + i) > 0.f;
        if (!any) {
            const int x = (int)(i % brick.v[0]);
            const size_t r = i / brick.v[0];
            const int y = (int)(r % brick.v[1]), z = (int)(r / brick.v[1]);
            psi[(x + poff.v[0]) + (size_t)pdim.v[0] * ((y + poff.v[1]) + (size_t)pdim.v[1] * (z + poff.v[2]))] = 0.f;
        }
    }
}

// Kernel image for one window size: zero-padded to the window and rotated so that the kernel centre
// (k / 2) sits at index 0 (FFTConvolution.java:506-516).  `out` must be zeroed; one thread per kernel voxel.
__global__ void __launch_bounds__(256) wrap_kernel_kernel(const float* __restrict__ k, Int3 kd, float* __restrict__ out, Int3 bd) {
    const int n = kd.v[0] * kd.v[1] * kd.v[2];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int x = i % kd.v[0], y = (i / kd.v[0]) % kd.v[1], z = i / (kd.v[0] * kd.v[1]);
        const int px = ((x - kd.v[0] / 2) % bd.v[0] + bd.v[0]) % bd.v[0], py = ((y - kd.v[1] / 2) % bd.v[1] + bd.v[1]) % bd.v[1],
                  pz = ((z - kd.v[2] / 2) % bd.v[2] + bd.v[2]) % bd.v[2];
        // windows are at least as long as the kernel, so no two kernel voxels land on the same slot
        out[px + (size_t)bd.v[0] * (py + (size_t)bd.v[1] * pz)] = k[i];
    }
}

cudaError_t launch_wrap_kernel(const float* d_kernel, const int k[3], float* d_out, const int b[3], cudaStream_t st) {
    Int3 kd, bd;
    for (int i = 0; i < 3; ++i) { kd.v[i] = k[i]; bd.v[i] = b[i]; }
    const int n = k[0] * k[1] * k[2];
    if (n <= 0) return cudaSuccess;
    wrap_kernel_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_kernel, kd, d_out, bd);
    return cudaGetLastError();
}

cudaError_t launch_mask_never_updated(const ViewPtrs& views, const int brick[3], float* psi, const int psi_dim[3], const int psi_off[3],
                                      cudaStream_t st) {
    const size_t n = (size_t)brick[0] * brick[1] * brick[2];
    if (n == 0) return cudaSuccess;
    const int blocks = (int)((n + 255) / 256 < (size_t)(sm_count() * 8) ? (n + 255) / 256 : (size_t)(sm_count() * 8));
    Int3 b, d, o;
    for (int i = 0; i < 3; ++i) { b.v[i] = brick[i]; d.v[i] = psi_dim[i]; o.v[i] = psi_off[i]; }
    mask_kernel<<<blocks, 256, 0, st>>>(views, b, psi, d, o);
    return cudaGetLastError();
}

// ---- neighbour-only barrier in peer memory ---------------------------------------------------------------
// phase 0 = signal, 1 = wait, 2 = both (one launch per exchange when every brick has its own GPU)
__global__ void peer_barrier_kernel(PeerFlags pf, unsigned int* my_flags, unsigned int value, unsigned int* status, int phase) {
    const int i = threadIdx.x;
    if (i >= pf.n) return;
    if (phase != 1) {
        // everything this stream wrote before (the step kernels' ghost voxels in the neighbours' arrays) is ordered before the flag
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pf.theirs[i] + pf.their_slot[i]), "r"(value) : "memory");
    }
    if (phase != 0) {
        const unsigned int* mine = my_flags + pf.my_slot[i];
        const long long t0 = clock64();
        for (;;) {
            unsigned int seen;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
            if ((int)(seen - value) >= 0) break;  // wrap-safe comparison
            if (clock64() - t0 > 20000000000LL) {  // ~10 s: a neighbour died or the callers' step sequences differ
                atomicExch(status, 1u);
                break;
            }
            __nanosleep(200);
        }
    }
}

cudaError_t launch_peer_barrier(const PeerFlags& pf, unsigned int* my_flags, unsigned int value, unsigned int* status, int phase, cudaStream_t st) {
    if (pf.n <= 0) return cudaSuccess;
    peer_barrier_kernel<<<1, 32, 0, st>>>(pf, my_flags, value, status, phase);
    return cudaGetLastError();
}

}  // namespace mvd
