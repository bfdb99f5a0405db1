// Register-resident FFT building blocks for sm_100a.
//
// Every line FFT is a two- or three-pass decimation: a thread keeps EPT points of its line in
// registers, runs one in-register radix-R butterfly per pass, multiplies by the inter-pass
// twiddles and exchanges through shared memory between passes.  Forward sign is -1 and nothing
// is scaled, matching the convention of the reference's FFT (J/fftMethods/FFTMethods.java:953
// forward -1 unscaled, :1021-1044 inverse +1); the 1/n scaling is folded into the kernel spectrum.
//
// All register arrays are indexed with compile-time constants only (no local memory).
#pragma once
#include <cuda_runtime.h>

namespace mvd {

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
// a * conj(b)
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {
    return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }
// multiply by -i (forward) or +i (inverse)
template <bool INV> __device__ __forceinline__ float2 mul_mi(float2 a) {
    return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
}
// a * w (forward) or a * conj(w) (inverse); w is always the forward twiddle exp(-2 pi i ...)
template <bool INV> __device__ __forceinline__ float2 ctw(float2 a, float2 w) { return INV ? cmulc(a, w) : cmul(a, w); }

template <bool INV> __device__ __forceinline__ void dft2(float2& a, float2& b) {
    const float2 t = a;
    a = cadd(t, b);
    b = csub(t, b);
}

// natural order in, natural order out
template <bool INV> __device__ __forceinline__ void dft4(float2& x0, float2& x1, float2& x2, float2& x3) {
    const float2 t0 = cadd(x0, x2), t1 = csub(x0, x2), t2 = cadd(x1, x3), t3 = mul_mi<INV>(csub(x1, x3));
    x0 = cadd(t0, t2);
    x1 = cadd(t1, t3);
    x2 = csub(t0, t2);
    x3 = csub(t1, t3);
}

template <bool INV> __device__ __forceinline__ void dft8(float2* x) {
    // even / odd radix-4, then combine with w8^k
    dft4<INV>(x[0], x[2], x[4], x[6]);
    dft4<INV>(x[1], x[3], x[5], x[7]);
    const float h = 0.70710678118654752440f;
    const float2 o1 = x[3], o2 = x[5], o3 = x[7];
    // w8^1 = (h, -+h), w8^2 = -+i, w8^3 = (-h, -+h)
    const float2 w1 = INV ? make_float2(h * (o1.x - o1.y), h * (o1.x + o1.y)) : make_float2(h * (o1.x + o1.y), h * (o1.y - o1.x));
    const float2 w2 = mul_mi<INV>(o2);
    const float2 w3 = INV ? make_float2(-h * (o3.x + o3.y), h * (o3.x - o3.y)) : make_float2(h * (o3.y - o3.x), -h * (o3.x + o3.y));
    const float2 e0 = x[0], e1 = x[2], e2 = x[4], e3 = x[6], o0 = x[1];
    x[0] = cadd(e0, o0); x[4] = csub(e0, o0);
    x[1] = cadd(e1, w1); x[5] = csub(e1, w1);
    x[2] = cadd(e2, w2); x[6] = csub(e2, w2);
    x[3] = cadd(e3, w3); x[7] = csub(e3, w3);
}

template <bool INV> __device__ __forceinline__ void dft16(float2* x) {
    // n = 4a + b ; k = c + 4d
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f, h = 0.70710678118654752440f;
#pragma unroll
    for (int b = 0; b < 4; ++b) dft4<INV>(x[b], x[4 + b], x[8 + b], x[12 + b]);  // over a -> index c at x[4c+b]
    // twiddle w16^(b*c): value at x[4c + b]
    const float sg = INV ? 1.f : -1.f;
    const float2 w1 = make_float2(c1, sg * s1), w2 = make_float2(h, sg * h), w3 = make_float2(s1, sg * c1);
    const float2 w6 = make_float2(-h, sg * h), w9 = make_float2(-c1, -sg * s1);
    x[4 * 1 + 1] = cmul(x[4 * 1 + 1], w1);
    x[4 * 1 + 2] = cmul(x[4 * 1 + 2], w2);
    x[4 * 1 + 3] = cmul(x[4 * 1 + 3], w3);
    x[4 * 2 + 1] = cmul(x[4 * 2 + 1], w2);
    x[4 * 2 + 2] = mul_mi<INV>(x[4 * 2 + 2]);  // w16^4 = -+i
    x[4 * 2 + 3] = cmul(x[4 * 2 + 3], w6);
    x[4 * 3 + 1] = cmul(x[4 * 3 + 1], w3);
    x[4 * 3 + 2] = cmul(x[4 * 3 + 2], w6);
    x[4 * 3 + 3] = cmul(x[4 * 3 + 3], w9);
#pragma unroll
    for (int c = 0; c < 4; ++c) dft4<INV>(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);  // over b -> d at x[4c+d]
    // x[4c + d] holds X[c + 4d]; transpose to natural order
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int d = c + 1; d < 4; ++d) {
            const float2 t = x[4 * c + d];
            x[4 * c + d] = x[4 * d + c];
            x[4 * d + c] = t;
        }
}

template <int R, bool INV> __device__ __forceinline__ void dftR(float2* x) {
    static_assert(R == 2 || R == 4 || R == 8 || R == 16, "unsupported radix");
    if constexpr (R == 2) dft2<INV>(x[0], x[1]);
    else if constexpr (R == 4) dft4<INV>(x[0], x[1], x[2], x[3]);
    else if constexpr (R == 8) dft8<INV>(x);
    else dft16<INV>(x);
}

// twiddle tables in global memory: tw_N[n] = exp(-2*pi*i*n/N), filled by the host in double precision
struct Twiddles {
    const float2* t[11];  // index log2(N), N up to 1024
};

template <int N> __device__ __forceinline__ constexpr int ilog2() {
    int l = 0;
    for (int n = N; n > 1; n >>= 1) ++l;
    return l;
}

// ---------------------------------------------------------------------------------------------
// Line FFT of N = EPT * T points owned by T threads, thread t holding v[m] = x[t + m*T].
// Result in the same ownership: v[m] = X[t + m*T].
//
// Passes (Stockham autosort, radices R0 = EPT first, then up to two more):  after a pass with
// accumulated span NS the element j = t + b*T (b-th butterfly of the thread) has k = j % NS and
// its R outputs go to positions (j / NS) * NS * R + k + r * NS.  Inter-pass twiddles
// w_N^(k * r * N/(NS*R)) are read from a table `tw` of N entries (anything indexable: a shared-memory
// pointer or the constant bank) -- one LDS per
// twiddle instead of a chain of complex multiplications.
//
// put(idx, val) / get(idx) address the line's exchange storage (any layout), sync() orders the
// exchange among the line's threads.
template <int N, int EPT, int R, int NS, bool INV, bool LAST, class Tw, class Put>
__device__ __forceinline__ void fft_pass(float2 (&v)[EPT], int t, const Tw& tw, Put put) {
    constexpr int T = N / EPT, NB = EPT / R;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        const int j = t + b * T;
        const int k = j & (NS - 1);
        float2 x[R];
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] = v[b + r * NB];
        if constexpr (NS > 1) {
            constexpr int STEP = N / (NS * R);
            const int base = k * STEP;  // < N / R
#pragma unroll
            for (int r = 1; r < R; ++r) x[r] = ctw<INV>(x[r], tw[base * r]);
        }
        dftR<R, INV>(x);
        if constexpr (LAST) {
#pragma unroll
            for (int r = 0; r < R; ++r) v[b + r * NB] = x[r];
        } else {
            const int j0 = (j - k) * R + k;
#pragma unroll
            for (int r = 0; r < R; ++r) put(j0 + r * NS, x[r]);
        }
    }
}

template <int N, int EPT, bool INV, class Tw, class Put, class Get, class Sync>
__device__ __forceinline__ void line_fft(float2 (&v)[EPT], int t, const Tw& tw, Put put, Get get, Sync sync) {
    constexpr int T = N / EPT;
    static_assert(N % EPT == 0 && N >= EPT, "line too short for EPT");
    constexpr int R0 = EPT;
    constexpr int REM1 = N / R0;
    if constexpr (REM1 == 1) {
        fft_pass<N, EPT, R0, 1, INV, true>(v, t, tw, put);
    } else {
        fft_pass<N, EPT, R0, 1, INV, false>(v, t, tw, put);
        sync();
#pragma unroll
        for (int m = 0; m < EPT; ++m) v[m] = get(t + m * T);
        constexpr int R1 = REM1 >= EPT ? EPT : REM1;
        constexpr int REM2 = REM1 / R1;
        if constexpr (REM2 == 1) {
            fft_pass<N, EPT, R1, R0, INV, true>(v, t, tw, put);
        } else {
            sync();
            fft_pass<N, EPT, R1, R0, INV, false>(v, t, tw, put);
            sync();
#pragma unroll
            for (int m = 0; m < EPT; ++m) v[m] = get(t + m * T);
            constexpr int R2 = REM2;
            static_assert(R2 <= EPT, "at most three passes supported");
            fft_pass<N, EPT, R2, R0 * R1, INV, true>(v, t, tw, put);
        }
    }
}

}  // namespace mvd
