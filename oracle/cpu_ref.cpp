// CPU restatement in C++17 (std::thread; this image's gcc has no OpenMP runtime) of the reference's FFT convolution and of the deconvolution loop built on
// it -- TEST INFRASTRUCTURE ONLY (second, independent oracle next to oracle_np.py; see oracle/README.md).
//
// Unlike oracle_np.py (float64, scipy FFT on the explicitly extended array) this file follows the
// reference's own route, in float32:
//   FFTConvolution.convolve      J/fftMethods/FFTConvolution.java:470-551 (vendored copy of the imglib2 class the
//                                hot path imports, MVDeconFFT.java:11): extend the image interval to
//                                img + K - 1, pad to a transform-friendly size centred with the odd pixel on the
//                                right (FFTMethods.java:1114-1137), read the padded interval from the EXTENDED image
//                                accessor (mirror-single / constant / zero), zero-pad the kernel to the same size and
//                                view it periodically from its centre K/2 (:506-516), R2C along x then C2C along y, z
//                                (FFTMethods.java:565-730, forward sign -1 unscaled :953), optional complex conjugate
//                                of the kernel spectrum (:95, :528-531), pointwise product (:607-614), inverse C2C
//                                z, y and C2R x scaled by 1/n per dimension (:1021-1091), crop (:1152-1175).
//   transform sizes              Mines JTK (edu.mines.jtk.dsp.FftComplex / FftReal, version managed by pom-fiji,
//                                not in the tree): prime-factor FFT, valid lengths are products of mutually prime
//                                factors from {2,3,4,5,7,8,9,11,13,16}.  nfftFast picks among the valid lengths by a
//                                cost table that is not available here; nfft_small below returns the SMALLEST valid
//                                length >= n (JTK's nfftSmall).  The choice changes float rounding only.
//   loop                         MVDeconvolution.java:333-444 (runIteration), :473-525 (computeQuotient),
//                                :582-705 (computeFinalValues / computeNextValue / tikhonov), whole-image CPU
//                                branch of convolve1 / convolve2 (MVDeconFFT.java:389-413, 480-501).
//
// The FFT itself is a plain mixed-radix decimation-in-time transform written for this file (radices 2,3,4,5,7,11,13
// by generic small DFTs): not JTK's prime-factor codelets, so values agree with the reference to float rounding only.
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <thread>

// static split of [0, n) over the host threads (REF_THREADS or hardware_concurrency)
static int g_threads = 0;
extern "C" int ref_num_threads() {
    if (g_threads <= 0) {
        const char* e = std::getenv("REF_THREADS");
        g_threads = e ? std::atoi(e) : (int)std::thread::hardware_concurrency();
        if (g_threads <= 0) g_threads = 1;
    }
    return g_threads;
}
extern "C" void ref_set_num_threads(int n) { g_threads = n > 0 ? n : 1; }
template <class F> static void parallel_for(long long n, F&& body) {  // body(begin, end)
    const int T = (int)std::min<long long>(ref_num_threads(), std::max<long long>(n, 1));
    if (T <= 1) { body(0LL, n); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t) {
        const long long lo = n * t / T, hi = n * (t + 1) / T;
        th.emplace_back([&body, lo, hi] { body(lo, hi); });
    }
    for (auto& x : th) x.join();
}

typedef std::complex<float> cf;

// ---- transform sizes ----------------------------------------------------------------------------------
static bool pfa_valid(int n) {
    // product of mutually prime factors from {2,3,4,5,7,8,9,11,13,16}: 2^a (a<=4) * 3^b (b<=2) * 5^c * 7^d * 11^e * 13^f, c,d,e,f <= 1
    const int primes[6] = {2, 3, 5, 7, 11, 13}, maxpow[6] = {4, 2, 1, 1, 1, 1};
    for (int i = 0; i < 6; ++i) {
        int c = 0;
        while (n % primes[i] == 0) { n /= primes[i]; ++c; }
        if (c > maxpow[i]) return false;
    }
    return n == 1;
}
extern "C" int ref_nfft_small_complex(int n) {
    for (int m = std::max(n, 1); m <= 720720; ++m)
        if (pfa_valid(m)) return m;
    return -1;
}
extern "C" int ref_nfft_small_real(int n) {  // FftReal: even length, half of it a valid complex length
    const int h = ref_nfft_small_complex((std::max(n, 2) + 1) / 2);
    return h < 0 ? -1 : 2 * h;
}

// ---- 1-D complex FFT, any length with small prime factors ---------------------------------------------
struct Plan {
    int n = 0;
    std::vector<int> factors;
    std::vector<cf> roots;  // exp(-2 pi i j / n)
};
static Plan make_plan(int n) {
    Plan p;
    p.n = n;
    int m = n;
    while (m % 4 == 0) { p.factors.push_back(4); m /= 4; }
    for (int f : {2, 3, 5, 7, 11, 13})
        while (m % f == 0) { p.factors.push_back(f); m /= f; }
    if (m != 1) p.factors.push_back(m);  // generic O(m^2) for anything else
    p.roots.resize(n);
    for (int j = 0; j < n; ++j) {
        const double a = -2.0 * M_PI * (double)j / (double)n;
        p.roots[j] = cf((float)std::cos(a), (float)std::sin(a));
    }
    return p;
}
static inline cf root(const Plan& p, long long idx, bool inverse) {
    const cf w = p.roots[(size_t)(idx % p.n)];
    return inverse ? std::conj(w) : w;
}
static void fft_rec(const Plan& p, const cf* in, cf* out, int n, int stride, size_t fi, bool inverse) {
    if (n == 1) { out[0] = in[0]; return; }
    const int r = p.factors[fi], m = n / r;
    for (int q = 0; q < r; ++q) fft_rec(p, in + (size_t)q * stride, out + (size_t)q * m, m, stride * r, fi + 1, inverse);
    cf t[16], y[16];
    std::vector<cf> tbig, ybig;
    cf* tp = t;
    cf* yp = y;
    if (r > 16) { tbig.resize(r); ybig.resize(r); tp = tbig.data(); yp = ybig.data(); }
    const long long step = p.n / n;  // w_n = w_N^step
    for (int k = 0; k < m; ++k) {
        for (int q = 0; q < r; ++q) tp[q] = out[(size_t)q * m + k] * root(p, step * (long long)q * k, inverse);
        for (int j = 0; j < r; ++j) {
            cf s = tp[0];
            for (int q = 1; q < r; ++q) s += tp[q] * root(p, (long long)(p.n / r) * ((long long)j * q % r), inverse);
            yp[j] = s;
        }
        for (int j = 0; j < r; ++j) out[(size_t)j * m + k] = yp[j];
    }
}
static void fft_line(const Plan& p, cf* x, cf* tmp, bool inverse) {
    fft_rec(p, x, tmp, p.n, 1, 0, inverse);
    std::memcpy(x, tmp, sizeof(cf) * p.n);
}

// ---- out-of-bounds rules of the image accessor ----------------------------------------------------------
enum { EXT_ZERO = 0, EXT_CONST = 1, EXT_MIRROR_SINGLE = 2 };
static inline long long mirror_single(long long i, long long n) {
    if (n == 1) return 0;
    const long long p = 2 * n - 2;
    i %= p;
    if (i < 0) i += p;
    return i >= n ? p - i : i;
}

// ---- FFTConvolution.convolve ------------------------------------------------------------------------------
// f: dims (x,y,z), x fastest.  k: kdims.  out: dims.  conjugate != 0 applies FFTMethods.complexConjugate to the
// kernel spectrum.  padded_out (optional, 3 ints) receives the transform sizes.
extern "C" int ref_fft_convolve(const float* f, const int* dims, const float* k, const int* kdims, int ext_mode, float ext_value,
                                int conjugate, float* out, int* padded_out) {
    long long pad[3], imin[3], kmin[3];
    for (int d = 0; d < 3; ++d) {
        const int nd = dims[d] + kdims[d] - 1;  // FFTConvolution.java:484-485
        pad[d] = d == 0 ? ref_nfft_small_real(nd) : ref_nfft_small_complex(nd);  // FFTMethods.java:1279-1291
        if (pad[d] < 0) return -1;
        imin[d] = -((pad[d] - dims[d]) / 2);   // paddingIntervalCentered: the odd pixel goes to the right
        kmin[d] = -((pad[d] - kdims[d]) / 2);
        if (padded_out) padded_out[d] = (int)pad[d];
    }
    const long long PX = pad[0], PY = pad[1], PZ = pad[2], HX = PX / 2 + 1;
    std::vector<cf> fi((size_t)HX * PY * PZ), fk((size_t)HX * PY * PZ);
    const Plan px = make_plan((int)PX), py = make_plan((int)PY), pz = make_plan((int)PZ);

    auto forward = [&](std::vector<cf>& spec, auto&& sample) {
        // real -> complex along x (FFTMethods.java:565-730)
        parallel_for(PZ * PY, [&](long long lo, long long hi) {
            std::vector<cf> line(PX), tmp(PX);
            for (long long i = lo; i < hi; ++i) {
                const long long z = i / PY, y = i % PY;
                for (long long x = 0; x < PX; ++x) line[x] = cf(sample(x, y, z), 0.f);
                fft_line(px, line.data(), tmp.data(), false);
                for (long long x = 0; x < HX; ++x) spec[(size_t)((z * PY + y) * HX + x)] = line[x];
            }
        });
        // complex -> complex along y, z (FFTMethods.java:797-930)
        parallel_for(PZ * HX, [&](long long lo, long long hi) {
            std::vector<cf> line(PY), tmp(PY);
            for (long long i = lo; i < hi; ++i) {
                const long long z = i / HX, x = i % HX;
                for (long long y = 0; y < PY; ++y) line[y] = spec[(size_t)((z * PY + y) * HX + x)];
                fft_line(py, line.data(), tmp.data(), false);
                for (long long y = 0; y < PY; ++y) spec[(size_t)((z * PY + y) * HX + x)] = line[y];
            }
        });
        parallel_for(PY * HX, [&](long long lo, long long hi) {
            std::vector<cf> line(PZ), tmp(PZ);
            for (long long i = lo; i < hi; ++i) {
                const long long y = i / HX, x = i % HX;
                for (long long z = 0; z < PZ; ++z) line[z] = spec[(size_t)((z * PY + y) * HX + x)];
                fft_line(pz, line.data(), tmp.data(), false);
                for (long long z = 0; z < PZ; ++z) spec[(size_t)((z * PY + y) * HX + x)] = line[z];
            }
        });
    };
    // image: the padded interval read from the extended accessor
    forward(fi, [&](long long x, long long y, long long z) -> float {
        long long c[3] = {x + imin[0], y + imin[1], z + imin[2]};
        for (int d = 0; d < 3; ++d) {
            if (c[d] < 0 || c[d] >= dims[d]) {
                if (ext_mode == EXT_MIRROR_SINGLE) c[d] = mirror_single(c[d], dims[d]);
                else return ext_mode == EXT_CONST ? ext_value : 0.f;
            }
        }
        return f[(size_t)((c[2] * dims[1] + c[1]) * dims[0] + c[0])];
    });
    // kernel: zero-padded to the same size, periodic, viewed from its centre (FFTConvolution.java:506-516)
    forward(fk, [&](long long x, long long y, long long z) -> float {
        long long c[3] = {x, y, z};
        for (int d = 0; d < 3; ++d) {
            // position min + j of the periodic view of the padded kernel interval [kmin, kmin + pad)
            long long q = kdims[d] / 2 + c[d];
            q = ((q - kmin[d]) % pad[d] + pad[d]) % pad[d] + kmin[d];
            if (q < 0 || q >= kdims[d]) return 0.f;
            c[d] = q;
        }
        return k[(size_t)((c[2] * kdims[1] + c[1]) * kdims[0] + c[0])];
    });
    const size_t ns = fi.size();
    parallel_for((long long)ns, [&](long long lo, long long hi) {
        for (long long i = lo; i < hi; ++i) fi[i] *= conjugate ? std::conj(fk[i]) : fk[i];  // :528-531, :607-614
    });
    // inverse: z, y complex, then complex -> real along x, each scaled by 1 / n (FFTMethods.java:934-1108)
    parallel_for(PY * HX, [&](long long lo, long long hi) {
        std::vector<cf> line(PZ), tmp(PZ);
        for (long long i = lo; i < hi; ++i) {
            const long long y = i / HX, x = i % HX;
            for (long long z = 0; z < PZ; ++z) line[z] = fi[(size_t)((z * PY + y) * HX + x)];
            fft_line(pz, line.data(), tmp.data(), true);
            for (long long z = 0; z < PZ; ++z) fi[(size_t)((z * PY + y) * HX + x)] = line[z] / (float)PZ;
        }
    });
    parallel_for(PZ * HX, [&](long long lo, long long hi) {
        std::vector<cf> line(PY), tmp(PY);
        for (long long i = lo; i < hi; ++i) {
            const long long z = i / HX, x = i % HX;
            for (long long y = 0; y < PY; ++y) line[y] = fi[(size_t)((z * PY + y) * HX + x)];
            fft_line(py, line.data(), tmp.data(), true);
            for (long long y = 0; y < PY; ++y) fi[(size_t)((z * PY + y) * HX + x)] = line[y] / (float)PY;
        }
    });
    // unpaddingIntervalCentered (FFTMethods.java:1152-1175): min = difference / 2 in every dimension
    const long long ox = (PX - dims[0]) / 2, oy = (PY - dims[1]) / 2, oz = (PZ - dims[2]) / 2;
    parallel_for((long long)dims[2] * dims[1], [&](long long lo, long long hi) {
        std::vector<cf> line(PX), tmp(PX);
        for (long long i = lo; i < hi; ++i) {
            const long long z = i / dims[1], y = i % dims[1];
            const cf* row = &fi[(size_t)(((z + oz) * PY + (y + oy)) * HX)];
            for (long long x = 0; x < HX; ++x) line[x] = row[x];
            for (long long x = HX; x < PX; ++x) line[x] = std::conj(row[PX - x]);
            fft_line(px, line.data(), tmp.data(), true);
            for (long long x = 0; x < dims[0]; ++x) out[(size_t)((z * dims[1] + y) * dims[0] + x)] = line[x + ox].real() / (float)PX;
        }
    });
    return 0;
}

// ---- elementwise passes ------------------------------------------------------------------------------------
// MVDeconvolution.java:473-525
extern "C" void ref_compute_quotient(float* psi_blurred, const float* img, long long n) {
    parallel_for(n, [&](long long lo, long long hi) {
        for (long long i = lo; i < hi; ++i) psi_blurred[i] = img[i] > 0.f ? img[i] / psi_blurred[i] : 1.f;
    });
}
static inline double tikhonov(double value, double lambda) { return (std::sqrt(1.0 + 2.0 * lambda * value) - 1.0) / lambda; }  // :705
// MVDeconvolution.java:582-703; weight == nullptr: constant weight_const.  Returns sum and max of |change|.
extern "C" void ref_compute_final_values(float* psi, const float* integral, const float* weight, float weight_const, double lambda,
                                         float min_value, long long n, double* sum_change, double* max_change) {
    const int T = ref_num_threads();
    std::vector<double> sums(T, 0.0);
    std::vector<float> maxs(T, -1.f);
    // one portion per thread, partial sums added in portion order (the reference: one RealSum per portion, :600-640)
    parallel_for(T, [&](long long plo, long long phi) {
        for (long long t = plo; t < phi; ++t) {
            double sum = 0.0;
            float mx = -1.f;
            for (long long i = n * t / T; i < n * (t + 1) / T; ++i) {
                const float last = psi[i];
                const float value = last * integral[i];
                float adjusted;
                if (value > 0.f) adjusted = lambda > 0.0 ? (float)tikhonov((double)value, lambda) : value;
                else adjusted = min_value;
                const float next = std::isnan(adjusted) ? min_value : std::max(min_value, adjusted);
                const float w = weight ? weight[i] : weight_const;
                // -ffp-contract=off (Makefile): statement-wise float rounding like the reference, no fused multiply-add
                const float nw = last + ((next - last) * w);
                const float change = std::fabs(nw - last);
                sum += (double)change;
                mx = std::max(mx, change);
                psi[i] = nw;
            }
            sums[t] = sum;
            maxs[t] = mx;
        }
    });
    double sum = 0.0;
    float mx = -1.f;
    for (int t = 0; t < T; ++t) { sum += sums[t]; mx = std::max(mx, maxs[t]); }
    *sum_change = sum;
    *max_change = (double)mx;
}

// ---- the loop (whole-image CPU branch) ------------------------------------------------------------------------
// imgs / weights / kernels: arrays of V pointers; weights[v] may be null (constant weight_consts[v]).
// stats: iterations * V * 2 doubles (sum, max).  psi is updated in place; the final mask is the caller's.
extern "C" int ref_run_iterations(const float* const* imgs, const float* const* weights, const float* weight_consts,
                                  const float* const* kernel1, const float* const* kernel2, const int* kdims1, const int* kdims2, int V,
                                  const int* dims, int iterations, double lambda, float min_value, int conjugate, float* psi, double* stats) {
    const long long n = (long long)dims[0] * dims[1] * dims[2];
    std::vector<float> t1((size_t)n), t2((size_t)n);
    for (int it = 0; it < iterations; ++it)
        for (int v = 0; v < V; ++v) {  // strictly in order, psi updated after every view (:353-441)
            if (ref_fft_convolve(psi, dims, kernel1[v], kdims1 + 3 * v, EXT_MIRROR_SINGLE, 0.f, conjugate, t1.data(), nullptr)) return -1;
            ref_compute_quotient(t1.data(), imgs[v], n);
            if (ref_fft_convolve(t1.data(), dims, kernel2[v], kdims2 + 3 * v, EXT_CONST, 1.f, conjugate, t2.data(), nullptr)) return -1;
            double s, m;
            ref_compute_final_values(psi, t2.data(), weights[v], weight_consts[v], lambda, min_value, n, &s, &m);
            stats[2 * ((size_t)it * V + v)] = s;
            stats[2 * ((size_t)it * V + v) + 1] = m;
        }
    return 0;
}

