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

template <bool INV> __device__ __forceinline__ void dft32(float2* x) {
    // even / odd radix-16, then X[k] = E[k] + w32^k O[k], X[k+16] = E[k] - w32^k O[k]
    float2 e[16], o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { e[i] = x[2 * i]; o[i] = x[2 * i + 1]; }
    dft16<INV>(e);
    dft16<INV>(o);
    const float sg = INV ? -1.f : 1.f;  // conjugate the forward twiddles for the inverse
    { const float2 t = o[0]; x[0] = cadd(e[0], t); x[16] = csub(e[0], t); }
    { const float2 t = cmul(o[1], make_float2(0.98078528040323043f, sg * -0.19509032201612825f)); x[1] = cadd(e[1], t); x[17] = csub(e[1], t); }
    { const float2 t = cmul(o[2], make_float2(0.92387953251128674f, sg * -0.38268343236508978f)); x[2] = cadd(e[2], t); x[18] = csub(e[2], t); }
    { const float2 t = cmul(o[3], make_float2(0.83146961230254524f, sg * -0.55557023301960218f)); x[3] = cadd(e[3], t); x[19] = csub(e[3], t); }
    { const float2 t = cmul(o[4], make_float2(0.70710678118654757f, sg * -0.70710678118654746f)); x[4] = cadd(e[4], t); x[20] = csub(e[4], t); }
    { const float2 t = cmul(o[5], make_float2(0.55557023301960229f, sg * -0.83146961230254524f)); x[5] = cadd(e[5], t); x[21] = csub(e[5], t); }
    { const float2 t = cmul(o[6], make_float2(0.38268343236508984f, sg * -0.92387953251128674f)); x[6] = cadd(e[6], t); x[22] = csub(e[6], t); }
    { const float2 t = cmul(o[7], make_float2(0.19509032201612833f, sg * -0.98078528040323043f)); x[7] = cadd(e[7], t); x[23] = csub(e[7], t); }
    { const float2 t = mul_mi<INV>(o[8]); x[8] = cadd(e[8], t); x[24] = csub(e[8], t); }
    { const float2 t = cmul(o[9], make_float2(-0.19509032201612819f, sg * -0.98078528040323043f)); x[9] = cadd(e[9], t); x[25] = csub(e[9], t); }
    { const float2 t = cmul(o[10], make_float2(-0.38268343236508973f, sg * -0.92387953251128674f)); x[10] = cadd(e[10], t); x[26] = csub(e[10], t); }
    { const float2 t = cmul(o[11], make_float2(-0.55557023301960196f, sg * -0.83146961230254546f)); x[11] = cadd(e[11], t); x[27] = csub(e[11], t); }
    { const float2 t = cmul(o[12], make_float2(-0.70710678118654746f, sg * -0.70710678118654757f)); x[12] = cadd(e[12], t); x[28] = csub(e[12], t); }
    { const float2 t = cmul(o[13], make_float2(-0.83146961230254535f, sg * -0.55557023301960218f)); x[13] = cadd(e[13], t); x[29] = csub(e[13], t); }
    { const float2 t = cmul(o[14], make_float2(-0.92387953251128674f, sg * -0.38268343236508989f)); x[14] = cadd(e[14], t); x[30] = csub(e[14], t); }
    { const float2 t = cmul(o[15], make_float2(-0.98078528040323043f, sg * -0.19509032201612861f)); x[15] = cadd(e[15], t); x[31] = csub(e[15], t); }
}

template <int R, bool INV> __device__ __forceinline__ void dftR(float2* x) {
    static_assert(R == 2 || R == 4 || R == 8 || R == 16 || R == 32, "unsupported radix");
    if constexpr (R == 2) dft2<INV>(x[0], x[1]);
    else if constexpr (R == 4) dft4<INV>(x[0], x[1], x[2], x[3]);
    else if constexpr (R == 8) dft8<INV>(x);
    else if constexpr (R == 16) dft16<INV>(x);
    else dft32<INV>(x);
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
