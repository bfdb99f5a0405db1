// Fast row passes of the xy kernels for 128-wide windows (H = 64 packed complex per row).
//
// Four threads share a row (q = lane & 3), eight rows share a warp, each thread keeps 16 points.
// The 64-point transform is two passes of radix 8 (two butterflies per thread and pass) with ONE
// exchange through the row's own 64 slots of the shared-memory plane, and the real<->half-complex
// untangle is done inside the thread: in the second pass a thread transforms the spectrum classes
// k mod 8 = {q, 8-q} ({0, 4} for q = 0), which hold both members of every conjugate pair
// (k, 64-k).  The input side owns adjacent points (n mod 8 = {2q, 2q+1}), so global accesses are
// float4.  Per element that is 3 shared-memory accesses (exchange write + read, natural write or
// read) where the generic pass needs 5, and none of them conflicts:
//
//   exchange slot of (class c, index i) = 8c + 4(i & 1) + ((i >> 1) ^ (c & 3))
//
// -- for a fixed c the four lanes cover one aligned quad of slots, for a fixed i the classes of
// the four lanes land in different quads, and the four rows of a half-warp sit 4 slots apart
// (pitch 68), so every 64-bit access touches the 16 slots of the 32 banks exactly once.
//
// Table layout (shared memory, filled once per CTA):
//   T1[(k1*4 + q)*2 + s]  forward: w64^((2q+s) k1)         inverse: w64^(class_s(q) k1)
//   TU[j*4 + q]           w128^(kp(q, j)), kp = k of the j-th conjugate pair of the thread
#pragma once
#include "fft_device.cuh"

namespace mvd {

__device__ __forceinline__ int row128_class_a(int q) { return q; }
__device__ __forceinline__ int row128_class_b(int q) { return q == 0 ? 4 : 8 - q; }
// k of pair j of thread q: q + 8j, except q = 0: {32, 8, 16, 24, 4, 12, 20, 28}
__device__ __forceinline__ int row128_pair_k(int q, int j) {
    if (q != 0) return q + 8 * j;
    return j == 0 ? 32 : (j < 4 ? 8 * j : 8 * j - 28);
}

// fills the 64 + 32 table entries; tw64 / tw128 are the global twiddle tables of length 64 / 128
__device__ __forceinline__ void row128_fill_tables(float2* T1, float2* TU, const float2* __restrict__ tw64,
                                                   const float2* __restrict__ tw128, bool inverse, int tid, int nthreads) {
    for (int i = tid; i < 64; i += nthreads) {
        const int s = i & 1, q = (i >> 1) & 3, k1 = i >> 3;
        const int c = inverse ? (s ? row128_class_b(q) : row128_class_a(q)) : 2 * q + s;
        T1[i] = __ldg(&tw64[(c * k1) & 63]);
    }
    for (int i = tid; i < 32; i += nthreads) {
        const int q = i & 3, j = i >> 2;
        TU[i] = __ldg(&tw128[row128_pair_k(q, j)]);
    }
}

struct Row128Ptrs {
    float2 *w[4];   // rp + (q ^ x)
    float2 *ra[4];  // rp + ((8 kA + (kA & 3)) ^ x)
    float2 *rb[4];  // rp + ((8 kB + (kB & 3)) ^ x)
};
__device__ __forceinline__ Row128Ptrs row128_ptrs(float2* rp, int q) {
    Row128Ptrs p;
    const int kA = row128_class_a(q), kB = row128_class_b(q);
    const int cA = 8 * kA + (kA & 3), cB = 8 * kB + (kB & 3);
#pragma unroll
    for (int x = 0; x < 4; ++x) {
        p.w[x] = rp + (q ^ x);
        p.ra[x] = rp + (cA ^ x);
        p.rb[x] = rp + (cB ^ x);
    }
    return p;
}

// Forward: a[n1] = z[8 n1 + 2q], b[n1] = z[8 n1 + 2q + 1] (z[j] = (x[2j], x[2j+1]) of a real row).
// Leaves 2 * X[k], k = 0..63, in natural order in rp[0..63] (slot 0 packs X[0] and X[64]).
__device__ __forceinline__ void row128_forward(float2 (&a)[8], float2 (&b)[8], float2* rp, const float2* T1, const float2* TU, int q,
                                               unsigned mask) {
    dft8<false>(a);
    dft8<false>(b);
    const float4* t4 = reinterpret_cast<const float4*>(T1) + q;
#pragma unroll
    for (int k1 = 1; k1 < 8; ++k1) {
        const float4 w = t4[4 * k1];
        a[k1] = cmul(a[k1], make_float2(w.x, w.y));
        b[k1] = cmul(b[k1], make_float2(w.z, w.w));
    }
    const Row128Ptrs p = row128_ptrs(rp, q);
#pragma unroll
    for (int k1 = 0; k1 < 8; ++k1) {
        p.w[k1 & 3][8 * k1] = a[k1];
        p.w[k1 & 3][8 * k1 + 4] = b[k1];
    }
    __syncwarp(mask);
    float2 c[8], d[8];
#pragma unroll
    for (int n2 = 0; n2 < 8; ++n2) {
        c[n2] = p.ra[n2 >> 1][4 * (n2 & 1)];
        d[n2] = p.rb[n2 >> 1][4 * (n2 & 1)];
    }
    dft8<false>(c);  // c[k2] = Z[kA + 8 k2]
    dft8<false>(d);  // d[k2] = Z[kB + 8 k2]
    __syncwarp(mask);
    const bool q0 = q == 0;
    // conjugate pairs (P[j], Q[j]) = (Z[k], Z[64 - k]), k = kp(q, j)
    float2 P[8], Q[8];
    P[0] = q0 ? c[4] : c[0];
    P[1] = c[1]; P[2] = c[2]; P[3] = c[3];
    P[4] = q0 ? d[0] : c[4]; P[5] = q0 ? d[1] : c[5]; P[6] = q0 ? d[2] : c[6]; P[7] = q0 ? d[3] : c[7];
    Q[0] = q0 ? c[4] : d[7]; Q[1] = q0 ? c[7] : d[6]; Q[2] = q0 ? c[6] : d[5]; Q[3] = q0 ? c[5] : d[4];
    Q[4] = q0 ? d[7] : d[3]; Q[5] = q0 ? d[6] : d[2]; Q[6] = q0 ? d[5] : d[1]; Q[7] = q0 ? d[4] : d[0];
    // natural slots: k = q + 8j + adj(j)
    float2* const pk0 = rp + q + (q0 ? 32 : 0);
    float2* const pk13 = rp + q;
    float2* const pk47 = rp + q - (q0 ? 28 : 0);
    float2* const pm0 = rp + 64 - q - (q0 ? 32 : 0);
    float2* const pm13 = rp + 64 - q;
    float2* const pm47 = rp + 64 - q + (q0 ? 28 : 0);
    const float2* tu = TU + q;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        // X[k] (x2) = (Z[k] + conj Z[H-k]) - i w^k (Z[k] - conj Z[H-k]);  X[H-k] = conj(s) - i conj(w d)
        const float2 zp = cconj(Q[j]);
        const float2 s = cadd(P[j], zp), dd = csub(P[j], zp);
        const float2 wd = cmul(tu[4 * j], dd);
        const float2 xk = make_float2(s.x + wd.y, s.y - wd.x);
        const float2 xm = make_float2(s.x - wd.y, -s.y - wd.x);
        float2* pk = j == 0 ? pk0 : (j < 4 ? pk13 : pk47);
        float2* pm = j == 0 ? pm0 : (j < 4 ? pm13 : pm47);
        pk[8 * j] = xk;
        pm[-8 * j] = xm;
    }
    if (q0) rp[0] = make_float2(2.f * (c[0].x + c[0].y), 2.f * (c[0].x - c[0].y));
}

// Inverse: reads X[k] in natural order from rp[0..63]; returns e[k2] = z[2q + 8 k2], f[k2] = z[2q + 1 + 8 k2],
// i.e. (e[k2].x, e[k2].y, f[k2].x, f[k2].y) are the real samples x = 16 k2 + 4q .. + 3 (unnormalised).
__device__ __forceinline__ void row128_inverse(float2 (&e)[8], float2 (&f)[8], float2* rp, const float2* T1, const float2* TU, int q,
                                               unsigned mask) {
    const bool q0 = q == 0;
    const float2* const pk0 = rp + q + (q0 ? 32 : 0);
    const float2* const pk13 = rp + q;
    const float2* const pk47 = rp + q - (q0 ? 28 : 0);
    const float2* const pm0 = rp + 64 - q - (q0 ? 32 : 0);
    const float2* const pm13 = rp + 64 - q;
    const float2* const pm47 = rp + 64 - q + (q0 ? 28 : 0);
    const float2* tu = TU + q;
    float2 P[8], Q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float2* pk = j == 0 ? pk0 : (j < 4 ? pk13 : pk47);
        const float2* pm = j == 0 ? pm0 : (j < 4 ? pm13 : pm47);
        const float2 xk = pk[8 * j];
        const float2 xp = cconj(pm[-8 * j]);
        // Z[k] = (X[k] + conj X[H-k]) + i conj(w^k) (X[k] - conj X[H-k]);  Z[H-k] = conj(s) + i conj(wd)
        const float2 s = cadd(xk, xp), dd = csub(xk, xp);
        const float2 wd = cmulc(dd, tu[4 * j]);
        P[j] = make_float2(s.x - wd.y, s.y + wd.x);
        Q[j] = make_float2(s.x + wd.y, wd.x - s.y);
    }
    const float2 x0 = rp[0];
    const float2 z0 = make_float2(x0.x + x0.y, x0.x - x0.y);
    __syncwarp(mask);
    float2 c[8], d[8];
    c[0] = q0 ? z0 : P[0];
    c[1] = P[1]; c[2] = P[2]; c[3] = P[3];
    c[4] = q0 ? P[0] : P[4]; c[5] = q0 ? Q[3] : P[5]; c[6] = q0 ? Q[2] : P[6]; c[7] = q0 ? Q[1] : P[7];
    d[0] = q0 ? P[4] : Q[7]; d[1] = q0 ? P[5] : Q[6]; d[2] = q0 ? P[6] : Q[5]; d[3] = q0 ? P[7] : Q[4];
    d[4] = q0 ? Q[7] : Q[3]; d[5] = q0 ? Q[6] : Q[2]; d[6] = q0 ? Q[5] : Q[1]; d[7] = q0 ? Q[4] : Q[0];
    dft8<true>(c);
    dft8<true>(d);
    const float4* t4 = reinterpret_cast<const float4*>(T1) + q;
#pragma unroll
    for (int k1 = 1; k1 < 8; ++k1) {
        const float4 w = t4[4 * k1];
        c[k1] = cmulc(c[k1], make_float2(w.x, w.y));
        d[k1] = cmulc(d[k1], make_float2(w.z, w.w));
    }
    const Row128Ptrs p = row128_ptrs(rp, q);
#pragma unroll
    for (int k1 = 0; k1 < 8; ++k1) {
        p.ra[k1 >> 1][4 * (k1 & 1)] = c[k1];
        p.rb[k1 >> 1][4 * (k1 & 1)] = d[k1];
    }
    __syncwarp(mask);
#pragma unroll
    for (int n2 = 0; n2 < 8; ++n2) {
        e[n2] = p.w[n2 & 3][8 * n2];
        f[n2] = p.w[n2 & 3][8 * n2 + 4];
    }
    dft8<true>(e);
    dft8<true>(f);
}

}  // namespace mvd
