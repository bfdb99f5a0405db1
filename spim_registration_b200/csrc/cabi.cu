// extern "C" surface of libmvdecon.so (include/mvdecon.h).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <array>
#include <cstring>
#include <algorithm>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/mvdecon.h"
#include "engine.h"
#include "input_kernels.h"
#include "sep_kernels.h"
#include "volume_kernels.h"

using namespace mvd;

// ---- error latch ---------------------------------------------------------------------------
static thread_local std::string t_err;
static thread_local int t_err_code = 0;
static std::mutex g_err_mutex;
static std::string g_err;  // last error of any thread (legacy void symbols run on Java worker threads)
static int g_err_code = 0;

static int latch(int code, const std::string& msg) {
    t_err = msg;
    t_err_code = code;
    {
        std::lock_guard<std::mutex> lk(g_err_mutex);
        g_err = msg;
        g_err_code = code;
    }
    std::fprintf(stderr, "[libmvdecon] error %d: %s\n", code, msg.c_str());
    return code;
}

namespace {
struct StreamGuard {
    cudaStream_t st = nullptr;
    StreamGuard() { MVD_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)); }
    ~StreamGuard() { if (st) cudaStreamDestroy(st); }
};
}  // namespace

template <class F> static int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const CudaError& e) {
        cudaGetLastError();  // clear the sticky flag of non-fatal errors
        return latch(2, e.what());
    } catch (const std::invalid_argument& e) {
        return latch(1, e.what());
    } catch (const std::bad_alloc&) {
        return latch(3, "out of host memory");
    } catch (const std::exception& e) {
        return latch(4, e.what());
    } catch (...) {
        return latch(5, "unknown failure");
    }
}

extern "C" int mvdGetLastError(char* buf, int n) {
    int code = t_err_code;
    std::string msg = t_err;
    {
        std::lock_guard<std::mutex> lk(g_err_mutex);
        if (code == 0) {  // nothing failed on this thread: report the last failure of any thread
            code = g_err_code;
            msg = g_err;
        }
        if (code == 0 || g_err == msg) {  // the process-wide latch holds the error just reported
            g_err_code = 0;
            g_err.clear();
        }
    }
    t_err_code = 0;
    t_err.clear();
    if (buf && n > 0) {
        std::strncpy(buf, msg.c_str(), (size_t)n - 1);
        buf[n - 1] = 0;
    }
    return code;
}

extern "C" const char* mvdVersion(void) { return "mvdecon-b200 0.1 sm_100a (tiled overlap-save FFT, fused epilogues)"; }
extern "C" long long mvdKernelLaunchCount(void) { return g_kernel_launches.load(); }
extern "C" int mvdSetTuning(const char* name, int value) {
    if (!name) return -1;
    if (std::strcmp(name, "z_tma") == 0) return g_use_z_tma.exchange(value);
    if (std::strcmp(name, "inv_ring") == 0) return g_use_inv_ring.exchange(value);
    if (std::strcmp(name, "z_warp") == 0) return g_use_z_warp.exchange(value);
    return -1;
}

// ---- context -------------------------------------------------------------------------------
// device allocation owned by its holder (a view that fails half way through mvdAddView leaks nothing)
struct DevMem {
    float* p = nullptr;
    DevMem() = default;
    DevMem(const DevMem&) = delete;
    DevMem& operator=(const DevMem&) = delete;
    DevMem(DevMem&& o) noexcept : p(o.p) { o.p = nullptr; }
    DevMem& operator=(DevMem&& o) noexcept {
        if (this != &o) { reset(); p = o.p; o.p = nullptr; }
        return *this;
    }
    ~DevMem() { reset(); }
    void reset() { if (p) cudaFree(p); p = nullptr; }
    void alloc(size_t bytes) { reset(); MVD_CUDA(cudaMalloc(&p, bytes)); }
    operator float*() const { return p; }
};

struct ViewData {
    DevMem img;
    DevMem weight;
    float weight_const = 1.f;
    // upload of img / weight on the context's copy stream; the first step that touches the view waits for it
    cudaEvent_t ready = nullptr;
    bool ready_pending = false;
    ViewData() = default;
    ViewData(ViewData&& o) noexcept { *this = std::move(o); }
    ViewData& operator=(ViewData&& o) noexcept {
        img = std::move(o.img); weight = std::move(o.weight); weight_const = o.weight_const;
        if (ready) cudaEventDestroy(ready);
        ready = o.ready; o.ready = nullptr; ready_pending = o.ready_pending;
        plan1 = o.plan1; plan2 = o.plan2; spec1 = std::move(o.spec1); spec2 = std::move(o.spec2);
        sep_kmax1 = o.sep_kmax1; sep_kmax2 = o.sep_kmax2; taps1 = o.taps1; taps2 = o.taps2;
        return *this;
    }
    ~ViewData() { if (ready) cudaEventDestroy(ready); }
    int plan1 = -1, plan2 = -1;
    std::unique_ptr<KernelSpectra> spec1, spec2;
    // direct path: kernel i is rank-1 separable and at most 31 wide (sep_kernels.cu)
    int sep_kmax1 = 0, sep_kmax2 = 0;  // 0 = not separable
    SepTaps taps1, taps2;
};

// K[z][y][x] ~ a[z] b[y] c[x]?  The factors are the marginal sums (they average the rounding noise of
// kernels that were themselves computed by convolutions); accepted when sum |K - a b c| <= 1e-6 sum |K|.
static int separate_kernel(const float* k, const int kd[3], SepTaps& taps) {
    const int kmax = sep_kmax_for(kd[0], kd[1], kd[2]);
    if (kmax == 0) return 0;
    std::vector<double> mx(kd[0], 0.0), my(kd[1], 0.0), mz(kd[2], 0.0);
    double S = 0.0, A = 0.0;
    for (int z = 0; z < kd[2]; ++z)
        for (int y = 0; y < kd[1]; ++y)
            for (int x = 0; x < kd[0]; ++x) {
                const double v = k[((size_t)z * kd[1] + y) * kd[0] + x];
                mx[x] += v; my[y] += v; mz[z] += v; S += v; A += std::fabs(v);
            }
    if (!(std::fabs(S) > 1e-30) || !std::isfinite(S)) return 0;
    double err = 0.0;
    for (int z = 0; z < kd[2]; ++z)
        for (int y = 0; y < kd[1]; ++y)
            for (int x = 0; x < kd[0]; ++x)
                err += std::fabs((double)k[((size_t)z * kd[1] + y) * kd[0] + x] - mx[x] * my[y] * mz[z] / (S * S));
    if (!(err <= 1e-6 * A)) return 0;
    std::vector<float> fx(kd[0]), fy(kd[1]), fz(kd[2]);
    for (int x = 0; x < kd[0]; ++x) fx[x] = (float)(mx[x] / S);
    for (int y = 0; y < kd[1]; ++y) fy[y] = (float)(my[y] / S);
    for (int z = 0; z < kd[2]; ++z) fz[z] = (float)mz[z];
    std::memset(&taps, 0, sizeof(taps));
    make_window_taps(fx.data(), kd[0], kmax, taps.x);
    make_window_taps(fy.data(), kd[1], kmax, taps.y);
    make_window_taps(fz.data(), kd[2], kmax, taps.z);
    return kmax;
}

struct mvdContext {
    int device = 0;
    Device* dev = nullptr;
    int gdim[3], boff[3], bdim[3], glo[3], ghi[3];
    int adim[3];  // ghosted array extent
    double lambda = 0.0;
    float min_value = 1e-4f;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t copy_stream = nullptr;  // uploads of mvdAddView (created on first use)
    float* psi = nullptr;
    float* quot = nullptr;
    std::vector<ViewData> views;
    std::vector<std::unique_ptr<ConvPlan>> plans;
    Scratch scratch;
    std::unique_ptr<Profiler> prof;
    double* d_sum = nullptr;
    unsigned int* d_max = nullptr;
    int stat_capacity = 0;
    double* d_fi_sum = nullptr;
    unsigned long long* d_fi_count = nullptr;
    // fused ghost exchange (mvdAttachPeer): host copies and device tables of the neighbours
    int ghost[3] = {0, 0, 0};
    std::vector<PeerTarget> peers_psi, peers_quot;
    PeerTarget* d_peers_psi = nullptr;
    PeerTarget* d_peers_quot = nullptr;
    std::vector<void*> ipc_mappings;
    // neighbour-only barrier in peer memory (mvdExportFlags / mvdAttachPeerFlags / mvdPeerSignal / mvdPeerWait)
    unsigned int* d_flags = nullptr;  // slots 0 .. 31: written by the neighbours; slot 32: status of the bounded spin
    PeerFlags peer_flags{};
    void ensure_flags() {
        if (d_flags) return;
        MVD_CUDA(cudaMalloc(&d_flags, 64 * sizeof(unsigned int)));
        MVD_CUDA(cudaMemset(d_flags, 0, 64 * sizeof(unsigned int)));
    }
    // direct separable path
    bool direct_enabled = true;
    float* sep_tmp = nullptr;
    size_t sep_tmp_cap = 0;
    bool direct_ok() const { return direct_enabled; }
    float* sep_scratch(const Volume& src) {
        const size_t n = sep_tmp_elems(src, bdim);
        if (n > sep_tmp_cap) {
            MVD_CUDA(cudaStreamSynchronize(stream));
            if (sep_tmp) cudaFree(sep_tmp);
            sep_tmp = nullptr;
            sep_tmp_cap = 0;
            MVD_CUDA(cudaMalloc(&sep_tmp, n * sizeof(float)));
            sep_tmp_cap = n;
        }
        return sep_tmp;
    }

    size_t brick_elems() const { return (size_t)bdim[0] * bdim[1] * bdim[2]; }
    size_t array_elems() const { return (size_t)adim[0] * adim[1] * adim[2]; }
    bool has_ghost() const {
        for (int d = 0; d < 3; ++d)
            if (glo[d] || ghi[d]) return true;
        return false;
    }
    void activate() const { MVD_CUDA(cudaSetDevice(device)); }
    int plan_for(const int k[3]) {
        for (size_t i = 0; i < plans.size(); ++i)
            if (same_plan(*plans[i], bdim, k)) return (int)i;
        plans.push_back(make_plan(bdim, k));
        if (!peers_psi.empty()) partition_for_peers(*plans.back(), peers_psi.data(), (int)peers_psi.size());
        return (int)plans.size() - 1;
    }
    void ensure_stats(int n) {
        if (n <= stat_capacity) return;
        int cap = std::max(n, std::max(256, stat_capacity * 2));
        double* ns = nullptr;
        unsigned int* nm = nullptr;
        MVD_CUDA(cudaMalloc(&ns, cap * sizeof(double)));
        MVD_CUDA(cudaMalloc(&nm, cap * sizeof(unsigned int)));
        MVD_CUDA(cudaMemsetAsync(ns, 0, cap * sizeof(double), stream));
        MVD_CUDA(cudaMemsetAsync(nm, 0, cap * sizeof(unsigned int), stream));
        if (d_sum) {
            MVD_CUDA(cudaMemcpyAsync(ns, d_sum, stat_capacity * sizeof(double), cudaMemcpyDeviceToDevice, stream));
            MVD_CUDA(cudaMemcpyAsync(nm, d_max, stat_capacity * sizeof(unsigned int), cudaMemcpyDeviceToDevice, stream));
            MVD_CUDA(cudaStreamSynchronize(stream));
            cudaFree(d_sum);
            cudaFree(d_max);
        }
        d_sum = ns;
        d_max = nm;
        stat_capacity = cap;
    }
    Volume ghosted_volume(const float* data, int ext, float ext_value) const {
        Volume v;
        v.data = data;
        for (int d = 0; d < 3; ++d) { v.dim[d] = adim[d]; v.off[d] = glo[d]; v.gmin[d] = boff[d]; v.gdim[d] = gdim[d]; }
        v.ext = ext;
        v.ext_value = ext_value;
        return v;
    }
    EpilogueArgs epilogue_base(float* out) const {
        EpilogueArgs ep;
        std::memset(&ep, 0, sizeof(ep));
        ep.out = out;
        ep.lambda = lambda;
        {
            const double a = 2.0 * lambda;  // make_tikhonov (device_common.cuh)
            ep.tik_hi = (float)a;
            ep.tik_lo = (float)(a - (double)ep.tik_hi);
        }
        ep.min_value = min_value;
        for (int d = 0; d < 3; ++d) { ep.out_dim[d] = adim[d]; ep.out_off[d] = glo[d]; ep.aux_dim[d] = bdim[d]; }
        if (out == psi) { ep.peers = d_peers_psi; ep.n_peers = (int)peers_psi.size(); }
        else if (out == quot) { ep.peers = d_peers_quot; ep.n_peers = (int)peers_quot.size(); }
        return ep;
    }
    ViewPtrs view_ptrs() const {
        ViewPtrs p;
        p.n = (int)views.size();
        for (int i = 0; i < p.n; ++i) p.img[i] = views[i].img;
        return p;
    }
};

static void require(bool ok, const char* msg) {
    if (!ok) throw std::invalid_argument(msg);
}

// Ghost widths and array extent of a brick: the ghost layers are clipped at the volume border, and
// in x the low ghost width and the row pitch are rounded up to multiples of 4 so that rows stay
// float4-aligned (the extra voxels are never consumed; exchanges address the ghost layers through
// mvdArrayInfo).
static void ghost_geometry(const int gdim[3], const int boff[3], const int bdim[3], const int ghost[3], int glo[3], int ghi[3], int adim[3]) {
    for (int d = 0; d < 3; ++d) {
        glo[d] = std::min(ghost[d], boff[d]);
        ghi[d] = std::min(ghost[d], gdim[d] - (boff[d] + bdim[d]));
        if (d == 0 && ghost[0] > 0) {  // only bricks with x neighbours are padded: a whole volume keeps its own extent
            if (glo[0] > 0) glo[0] = (glo[0] + 3) / 4 * 4;
            const int rem = (bdim[0] + glo[0] + ghi[0]) % 4;
            if (rem) ghi[0] += 4 - rem;
        }
        adim[d] = bdim[d] + glo[d] + ghi[d];
    }
}

// copy between a dense brick-sized array and the interior of a ghosted array
static void copy_brick(const mvdContext& c, float* ghosted, float* brick, bool to_ghosted, cudaMemcpyKind kind) {
    cudaMemcpy3DParms p;
    std::memset(&p, 0, sizeof(p));
    cudaPitchedPtr g = make_cudaPitchedPtr(ghosted, (size_t)c.adim[0] * sizeof(float), c.adim[0], c.adim[1]);
    cudaPitchedPtr b = make_cudaPitchedPtr(brick, (size_t)c.bdim[0] * sizeof(float), c.bdim[0], c.bdim[1]);
    const cudaPos gpos = make_cudaPos((size_t)c.glo[0] * sizeof(float), c.glo[1], c.glo[2]);
    const cudaPos zero = make_cudaPos(0, 0, 0);
    if (to_ghosted) { p.srcPtr = b; p.srcPos = zero; p.dstPtr = g; p.dstPos = gpos; }
    else { p.srcPtr = g; p.srcPos = gpos; p.dstPtr = b; p.dstPos = zero; }
    p.extent = make_cudaExtent((size_t)c.bdim[0] * sizeof(float), c.bdim[1], c.bdim[2]);
    p.kind = kind;
    MVD_CUDA(cudaMemcpy3DAsync(&p, c.stream));
}

extern "C" int mvdCreate(mvdContext** out, int device, const int globalDim[3], const int brickOffset[3], const int brickDim[3],
                         const int ghost[3], double lambda, float minValue) {
    return guarded([&] {
        require(out && globalDim, "mvdCreate: null argument");
        *out = nullptr;
        std::unique_ptr<mvdContext> c(new mvdContext);
        c->device = device;
        c->dev = &Device::get(device);
        for (int d = 0; d < 3; ++d) {
            c->gdim[d] = globalDim[d];
            c->boff[d] = brickOffset ? brickOffset[d] : 0;
            c->bdim[d] = brickDim ? brickDim[d] : globalDim[d];
            require(c->gdim[d] > 0 && c->bdim[d] > 0 && c->boff[d] >= 0 && c->boff[d] + c->bdim[d] <= c->gdim[d],
                    "mvdCreate: brick outside the volume");
            c->ghost[d] = ghost ? ghost[d] : 0;
            require(c->ghost[d] >= 0, "mvdCreate: negative ghost width");
        }
        ghost_geometry(c->gdim, c->boff, c->bdim, c->ghost, c->glo, c->ghi, c->adim);
        require((double)c->array_elems() < 2147483647.0 * 4.0, "mvdCreate: brick too large");
        c->lambda = lambda;
        c->min_value = minValue;
        MVD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        c->own_stream = true;
        MVD_CUDA(cudaMalloc(&c->psi, c->array_elems() * sizeof(float)));
        MVD_CUDA(cudaMalloc(&c->quot, c->array_elems() * sizeof(float)));
        MVD_CUDA(cudaMalloc(&c->d_fi_sum, sizeof(double)));
        MVD_CUDA(cudaMalloc(&c->d_fi_count, sizeof(unsigned long long)));
        MVD_CUDA(launch_fill(c->psi, c->array_elems(), 0.f, c->stream));
        MVD_CUDA(launch_fill(c->quot, c->array_elems(), 1.f, c->stream));
        g_kernel_launches += 2;
        *out = c.release();
    });
}

static void detach_peers(mvdContext* c) {
    if (c->stream) cudaStreamSynchronize(c->stream);
    c->peers_psi.clear();
    c->peers_quot.clear();
    for (auto& plan : c->plans) partition_for_peers(*plan, nullptr, 0);
    if (c->d_peers_psi) cudaFree(c->d_peers_psi);
    if (c->d_peers_quot) cudaFree(c->d_peers_quot);
    c->d_peers_psi = c->d_peers_quot = nullptr;
    for (void* p : c->ipc_mappings) cudaIpcCloseMemHandle(p);
    c->ipc_mappings.clear();
    c->peer_flags.n = 0;
}

extern "C" int mvdDestroy(mvdContext* c) {
    if (!c) return 0;
    return guarded([&] {
        cudaSetDevice(c->device);
        if (c->stream) cudaStreamSynchronize(c->stream);
        if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); c->copy_stream = nullptr; }
        c->views.clear();
        c->plans.clear();
        c->prof.reset();
        detach_peers(c);
        if (c->sep_tmp) cudaFree(c->sep_tmp);
        if (c->psi) cudaFree(c->psi);
        if (c->quot) cudaFree(c->quot);
        if (c->d_flags) cudaFree(c->d_flags);
        if (c->d_sum) cudaFree(c->d_sum);
        if (c->d_max) cudaFree(c->d_max);
        if (c->d_fi_sum) cudaFree(c->d_fi_sum);
        if (c->d_fi_count) cudaFree(c->d_fi_count);
        if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
        delete c;
    });
}

extern "C" int mvdSetStream(mvdContext* c, void* s) {
    return guarded([&] {
        require(c, "mvdSetStream: null context");
        c->activate();
        MVD_CUDA(cudaStreamSynchronize(c->stream));
        if (c->own_stream) cudaStreamDestroy(c->stream);
        c->stream = (cudaStream_t)s;
        c->own_stream = false;
    });
}

static void check_peer_status(mvdContext* c);

extern "C" int mvdSynchronize(mvdContext* c) {
    return guarded([&] {
        require(c, "mvdSynchronize: null context");
        c->activate();
        MVD_CUDA(cudaStreamSynchronize(c->stream));
        check_peer_status(c);
    });
}

// ---- views ---------------------------------------------------------------------------------
// `volDim` / `volOff` (both or neither): img and weight are x-fastest host volumes of extent volDim and the brick
// sits at volOff inside them (the single-process multi-device API uploads bricks straight from the caller's
// whole-volume arrays); otherwise they are dense brick-sized arrays.
static void add_view(mvdContext* c, const float* img, const float* weight, float weightConst, const float* kernel1, const int k1[3],
                     const float* kernel2, const int k2[3], int memKind, const int* volDim, const int* volOff, int* viewIndex) {
    require(c && img && kernel1 && kernel2 && k1 && k2, "mvdAddView: null argument");
    require((int)c->views.size() < kMaxViews, "mvdAddView: too many views");
    c->activate();
    for (int d = 0; d < 3; ++d) {
        const int need1 = k1[d] - 1 - k1[d] / 2, need2 = k2[d] - 1 - k2[d] / 2;
        const bool lo_nb = c->boff[d] > 0, hi_nb = c->boff[d] + c->bdim[d] < c->gdim[d];
        if (lo_nb) require(c->glo[d] >= std::max(k1[d] / 2, need1) && c->glo[d] >= std::max(k2[d] / 2, need2),
                           "mvdAddView: ghost layer narrower than the kernel halo");
        if (hi_nb) require(c->ghi[d] >= std::max(k1[d] / 2, need1) && c->ghi[d] >= std::max(k2[d] / 2, need2),
                           "mvdAddView: ghost layer narrower than the kernel halo");
    }
    require(memKind == MVD_MEM_HOST || memKind == MVD_MEM_DEVICE || memKind == MVD_MEM_HOST_IMG_DEVICE_WEIGHT || memKind == MVD_MEM_HOST_ASYNC ||
                memKind == MVD_MEM_HOST_ASYNC_IMG_DEVICE_WEIGHT,
            "mvdAddView: bad memKind");
    const bool async_img = memKind == MVD_MEM_HOST_ASYNC || memKind == MVD_MEM_HOST_ASYNC_IMG_DEVICE_WEIGHT;
    require(!volDim || memKind == MVD_MEM_HOST || memKind == MVD_MEM_HOST_ASYNC, "mvdAddView: volume slices are host arrays");
    const cudaMemcpyKind kind = memKind == MVD_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    const cudaMemcpyKind wkind = (memKind == MVD_MEM_HOST || memKind == MVD_MEM_HOST_ASYNC) ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    if (!c->copy_stream) MVD_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    ViewData v;
    const size_t bytes = c->brick_elems() * sizeof(float);
    auto upload = [&](float* dst, const float* src, cudaMemcpyKind k) {
        if (!volDim) {
            MVD_CUDA(cudaMemcpyAsync(dst, src, bytes, k, c->copy_stream));
            return;
        }
        cudaMemcpy3DParms p;
        std::memset(&p, 0, sizeof(p));
        p.srcPtr = make_cudaPitchedPtr(const_cast<float*>(src), (size_t)volDim[0] * sizeof(float), volDim[0], volDim[1]);
        p.srcPos = make_cudaPos((size_t)volOff[0] * sizeof(float), volOff[1], volOff[2]);
        p.dstPtr = make_cudaPitchedPtr(dst, (size_t)c->bdim[0] * sizeof(float), c->bdim[0], c->bdim[1]);
        p.dstPos = make_cudaPos(0, 0, 0);
        p.extent = make_cudaExtent((size_t)c->bdim[0] * sizeof(float), c->bdim[1], c->bdim[2]);
        p.kind = k;
        MVD_CUDA(cudaMemcpy3DAsync(&p, c->copy_stream));
    };
    v.img.alloc(bytes);
    upload(v.img, img, kind);
    if (weight) {
        v.weight.alloc(bytes);
        if (memKind == MVD_MEM_HOST_ASYNC_IMG_DEVICE_WEIGHT) {
            // device -> device on the context's own stream (ordered before every step) and complete on return: the caller may
            // free the weight array at once, while the image upload is still in flight on the copy stream
            MVD_CUDA(cudaMemcpyAsync(v.weight, weight, bytes, cudaMemcpyDeviceToDevice, c->stream));
            MVD_CUDA(cudaStreamSynchronize(c->stream));
        } else {
            upload(v.weight, weight, wkind);
        }
    }
    MVD_CUDA(cudaEventCreateWithFlags(&v.ready, cudaEventDisableTiming));
    MVD_CUDA(cudaEventRecord(v.ready, c->copy_stream));
    v.ready_pending = true;
    v.weight_const = weightConst;
    v.plan1 = c->plan_for(k1);
    v.plan2 = c->plan_for(k2);
    {   // on a stream of their own: the spectra are built while the image and the weights upload
        StreamGuard aux;
        v.spec1 = build_spectra(*c->dev, *c->plans[v.plan1], kernel1, aux.st);
        v.spec2 = build_spectra(*c->dev, *c->plans[v.plan2], kernel2, aux.st);
    }
    v.sep_kmax1 = separate_kernel(kernel1, k1, v.taps1);
    v.sep_kmax2 = separate_kernel(kernel2, k2, v.taps2);
    // MVD_MEM_HOST_ASYNC: the caller keeps img / weight alive until the next mvdSynchronize / mvdRunIterations returns, so the
    // upload of this view overlaps the caller's next calls (and the first steps of the loop); otherwise the buffers may go
    // away after return
    if (!async_img) MVD_CUDA(cudaStreamSynchronize(c->copy_stream));
    c->views.push_back(std::move(v));
    if (viewIndex) *viewIndex = (int)c->views.size() - 1;
}

extern "C" int mvdAddView(mvdContext* c, const float* img, const float* weight, float weightConst, const float* kernel1,
                          const int k1[3], const float* kernel2, const int k2[3], int memKind, int* viewIndex) {
    return guarded([&] { add_view(c, img, weight, weightConst, kernel1, k1, kernel2, k2, memKind, nullptr, nullptr, viewIndex); });
}

// every kernel that reads a view's image or weights runs behind its upload
static void wait_view(mvdContext* c, ViewData& vd) {
    if (vd.ready_pending) {
        MVD_CUDA(cudaStreamWaitEvent(c->stream, vd.ready, 0));
        vd.ready_pending = false;
    }
}
static void wait_all_views(mvdContext* c) {
    for (auto& v : c->views) wait_view(c, v);
}

extern "C" int mvdNumViews(mvdContext* c) { return c ? (int)c->views.size() : -1; }

// ---- psi -----------------------------------------------------------------------------------
extern "C" int mvdFirstIterationSums(mvdContext* c, double* sum, long long* count) {
    return guarded([&] {
        require(c && sum && count, "mvdFirstIterationSums: null argument");
        require(!c->views.empty(), "mvdFirstIterationSums: no views");
        c->activate();
        MVD_CUDA(cudaMemsetAsync(c->d_fi_sum, 0, sizeof(double), c->stream));
        MVD_CUDA(cudaMemsetAsync(c->d_fi_count, 0, sizeof(unsigned long long), c->stream));
        wait_all_views(c);
        MVD_CUDA(launch_first_iteration_sums(c->view_ptrs(), c->brick_elems(), c->d_fi_sum, c->d_fi_count, c->stream));
        ++g_kernel_launches;
        unsigned long long cnt = 0;
        MVD_CUDA(cudaMemcpyAsync(sum, c->d_fi_sum, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        MVD_CUDA(cudaMemcpyAsync(&cnt, c->d_fi_count, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
        MVD_CUDA(cudaStreamSynchronize(c->stream));
        *count = (long long)cnt;
    });
}

extern "C" int mvdFillPsi(mvdContext* c, float value) {
    return guarded([&] {
        require(c, "mvdFillPsi: null context");
        c->activate();
        MVD_CUDA(launch_fill(c->psi, c->array_elems(), value, c->stream));
        ++g_kernel_launches;
    });
}

extern "C" int mvdInitPsiAverage(mvdContext* c, double* avg_out) {
    double sum = 0.0;
    long long count = 0;
    int rc = mvdFirstIterationSums(c, &sum, &count);
    if (rc) return rc;
    // MVDeconvolution.java:112-126: avg = sum / count, NaN -> 0.5
    double avg = sum / (double)count;
    if (std::isnan(avg)) avg = 0.5;
    if (avg_out) *avg_out = avg;
    return mvdFillPsi(c, (float)avg);
}

extern "C" int mvdSetPsi(mvdContext* c, const float* psi, int memKind) {
    return guarded([&] {
        require(c && psi, "mvdSetPsi: null argument");
        c->activate();
        copy_brick(*c, c->psi, const_cast<float*>(psi), true, memKind == MVD_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
        MVD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

extern "C" int mvdGetPsi(mvdContext* c, float* psi, int memKind) {
    return guarded([&] {
        require(c && psi, "mvdGetPsi: null argument");
        c->activate();
        copy_brick(*c, c->psi, psi, false, memKind == MVD_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost);
        MVD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

// ---- the loop ------------------------------------------------------------------------------
static void step_quotient(mvdContext* c, int v) {
    ViewData& vd = c->views[v];
    wait_view(c, vd);
    // conv1: psi extended by mirror-single (MVDeconFFTThreads.java:23,57), fused quotient (MVDeconvolution.java:473-525)
    const Volume src = c->ghosted_volume(c->psi, EXT_MIRROR_SINGLE, 0.f);
    EpilogueArgs ep = c->epilogue_base(c->quot);
    ep.img = vd.img;
    if (vd.sep_kmax1 && c->direct_ok()) {
        MVD_CUDA(launch_separable(src, c->bdim, vd.taps1, vd.sep_kmax1, EPI_QUOTIENT, ep, c->sep_scratch(src), c->stream));
        g_kernel_launches += 2;
        return;
    }
    run_convolution(*c->dev, *c->plans[vd.plan1], *vd.spec1, src, EPI_QUOTIENT, ep, c->scratch, c->stream, c->prof.get());
}

static void step_update(mvdContext* c, int v, int slot) {
    ViewData& vd = c->views[v];
    wait_view(c, vd);
    c->ensure_stats(slot + 1);
    MVD_CUDA(cudaMemsetAsync(c->d_sum + slot, 0, sizeof(double), c->stream));
    MVD_CUDA(cudaMemsetAsync(c->d_max + slot, 0, sizeof(unsigned int), c->stream));
    // conv2: quotient extended by the constant 1 (MVDeconFFTThreads.java:42,82), fused update (MVDeconvolution.java:582-703)
    const Volume src = c->ghosted_volume(c->quot, EXT_CONST, 1.f);
    EpilogueArgs ep = c->epilogue_base(c->psi);
    ep.weight = vd.weight;
    ep.weight_const = vd.weight_const;
    ep.sum_change = c->d_sum + slot;
    ep.max_change_bits = c->d_max + slot;
    if (vd.sep_kmax2 && c->direct_ok()) {
        MVD_CUDA(launch_separable(src, c->bdim, vd.taps2, vd.sep_kmax2, EPI_UPDATE, ep, c->sep_scratch(src), c->stream));
        g_kernel_launches += 2;
        return;
    }
    run_convolution(*c->dev, *c->plans[vd.plan2], *vd.spec2, src, EPI_UPDATE, ep, c->scratch, c->stream, c->prof.get());
}

static void read_stats(mvdContext* c, int first, int n, double* sum, double* mx) {
    check_peer_status(c);
    if (n <= 0) return;
    require(first >= 0 && first + n <= c->stat_capacity, "statistics slot out of range");
    std::vector<unsigned int> bits(n);
    std::vector<double> s(n);
    MVD_CUDA(cudaMemcpyAsync(s.data(), c->d_sum + first, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    MVD_CUDA(cudaMemcpyAsync(bits.data(), c->d_max + first, n * sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream));
    MVD_CUDA(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < n; ++i) {
        if (sum) sum[i] = s[i];
        if (mx) {
            float f;
            std::memcpy(&f, &bits[i], sizeof(f));
            mx[i] = (double)f;
        }
    }
}

extern "C" int mvdRunIterations(mvdContext* c, int numIterations, double* sumChange, double* maxChange) {
    return guarded([&] {
        require(c, "mvdRunIterations: null context");
        require(!c->views.empty(), "mvdRunIterations: no views");
        require(!c->has_ghost(), "mvdRunIterations: context has ghost layers, use the step API");
        require(numIterations >= 0, "mvdRunIterations: negative iteration count");
        c->activate();
        const int nv = (int)c->views.size();
        c->ensure_stats(numIterations * nv);
        for (int it = 0; it < numIterations; ++it)
            for (int v = 0; v < nv; ++v) {  // strictly in order, psi updated after every view (MVDeconvolution.java:353-441)
                step_quotient(c, v);
                step_update(c, v, it * nv + v);
            }
        if (sumChange || maxChange) read_stats(c, 0, numIterations * nv, sumChange, maxChange);
        else MVD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

extern "C" int mvdStepQuotient(mvdContext* c, int view) {
    return guarded([&] {
        require(c && view >= 0 && view < (int)c->views.size(), "mvdStepQuotient: bad view");
        c->activate();
        step_quotient(c, view);
    });
}

extern "C" int mvdStepUpdate(mvdContext* c, int view, int statSlot) {
    return guarded([&] {
        require(c && view >= 0 && view < (int)c->views.size() && statSlot >= 0, "mvdStepUpdate: bad argument");
        c->activate();
        step_update(c, view, statSlot);
    });
}

extern "C" int mvdGetStats(mvdContext* c, int firstSlot, int numSlots, double* sumChange, double* maxChange) {
    return guarded([&] {
        require(c, "mvdGetStats: null context");
        c->activate();
        read_stats(c, firstSlot, numSlots, sumChange, maxChange);
    });
}

extern "C" int mvdProfileStart(mvdContext* c) {
    return guarded([&] {
        require(c, "mvdProfileStart: null context");
        c->prof.reset(new Profiler);
    });
}

extern "C" int mvdProfileStop(mvdContext* c, double ms[3], long long launches[3], double bytes[3]) {
    return guarded([&] {
        require(c && c->prof && ms && launches && bytes, "mvdProfileStop: profiling not started");
        c->activate();
        c->prof->collect(ms, launches, bytes);
        c->prof.reset();
    });
}

extern "C" int mvdArrayInfo(mvdContext* c, int which, void** ptr, int dim[3], int off[3]) {
    return guarded([&] {
        require(c && (which == MVD_ARRAY_PSI || which == MVD_ARRAY_QUOTIENT), "mvdArrayInfo: bad argument");
        if (ptr) *ptr = which == MVD_ARRAY_PSI ? c->psi : c->quot;
        for (int d = 0; d < 3; ++d) {
            if (dim) dim[d] = c->adim[d];
            if (off) off[d] = c->glo[d];
        }
    });
}

static void convolve_device(Device& dev, const float* d_src, const int dim[3], const float* kernel, const int k[3], int ext,
                            float ext_value, float* d_out, cudaStream_t st);

// ---- direct separable path -------------------------------------------------------------------
extern "C" int mvdSetDirectPath(mvdContext* c, int enable) {
    return guarded([&] {
        require(c, "mvdSetDirectPath: null context");
        c->direct_enabled = enable != 0;
    });
}

extern "C" int mvdViewUsesDirectPath(mvdContext* c, int view, int* conv1, int* conv2) {
    return guarded([&] {
        require(c && view >= 0 && view < (int)c->views.size(), "mvdViewUsesDirectPath: bad view");
        if (conv1) *conv1 = c->views[view].sep_kmax1 && c->direct_ok() ? c->views[view].sep_kmax1 : 0;
        if (conv2) *conv2 = c->views[view].sep_kmax2 && c->direct_ok() ? c->views[view].sep_kmax2 : 0;
    });
}

// The reference's second native hook: in-place separable convolution of a whole image or block with
// one kernel of diameter N per axis (CUDASeparableConvolution.java:13-17, called by
// CUDASeparableConvolutionFunctions.gauss, :127-199).  N <= 31 runs the direct kernels, 63 and 127 the FFT
// engine axis by axis (127 taps along x or y use the 256-point windows).
static int separable_hook(int N, float* image, const float* kx, const float* ky, const float* kz, int w, int h, int d, int cx, int cy,
                          int cz, int oob, float oobv, int dev_id) {
    const int rc = guarded([&] {
        require(image && w > 0 && h > 0 && d > 0, "convolve_N: bad image");
        require(oob >= 0 && oob <= 2, "convolve_N: outofbounds must be 0 (zero), 1 (value) or 2 (extend last pixel)");
        require((!cx || kx) && (!cy || ky) && (!cz || kz), "convolve_N: null kernel for an enabled axis");
        int n_dev = 0;
        MVD_CUDA(cudaGetDeviceCount(&n_dev));
        require(dev_id >= 0 && dev_id < n_dev, "convolve_N: no such CUDA device (this library has no CPU path)");
        MVD_CUDA(cudaSetDevice(dev_id));
        Device& dev = Device::get(dev_id);
        const size_t n = (size_t)w * h * d;
        struct Bufs {
            float *a = nullptr, *b = nullptr, *t = nullptr;
            cudaStream_t st = nullptr;
            ~Bufs() { if (a) cudaFree(a); if (b) cudaFree(b); if (t) cudaFree(t); if (st) cudaStreamDestroy(st); }
        } B;
        MVD_CUDA(cudaStreamCreateWithFlags(&B.st, cudaStreamNonBlocking));
        MVD_CUDA(cudaMalloc(&B.a, n * sizeof(float)));
        MVD_CUDA(cudaMalloc(&B.b, n * sizeof(float)));
        MVD_CUDA(cudaMemcpyAsync(B.a, image, n * sizeof(float), cudaMemcpyHostToDevice, B.st));
        const int dim[3] = {w, h, d};
        const int ext = oob == 0 ? EXT_ZERO : (oob == 1 ? EXT_CONST : EXT_CLAMP);
        const float one = 1.f;
        if (N <= kSepMaxTaps) {
            Volume v;
            v.data = B.a;
            for (int i = 0; i < 3; ++i) { v.dim[i] = dim[i]; v.off[i] = 0; v.gmin[i] = 0; v.gdim[i] = dim[i]; }
            v.ext = ext;
            v.ext_value = oobv;
            SepTaps taps;
            std::memset(&taps, 0, sizeof(taps));
            const int kmax = sep_kmax_for(N, N, N);
            make_window_taps(cx ? kx : &one, cx ? N : 1, kmax, taps.x);
            make_window_taps(cy ? ky : &one, cy ? N : 1, kmax, taps.y);
            make_window_taps(cz ? kz : &one, cz ? N : 1, kmax, taps.z);
            taps.per_pass_fill = 1;
            EpilogueArgs ep;
            std::memset(&ep, 0, sizeof(ep));
            ep.out = B.b;
            for (int i = 0; i < 3; ++i) { ep.out_dim[i] = dim[i]; ep.aux_dim[i] = dim[i]; }
            MVD_CUDA(cudaMalloc(&B.t, sep_tmp_elems(v, dim) * sizeof(float)));
            MVD_CUDA(launch_separable(v, dim, taps, kmax, EPI_STORE, ep, B.t, B.st));
            g_kernel_launches += 2;
            MVD_CUDA(cudaStreamSynchronize(B.st));  // the caller's array is only touched once everything has succeeded
            MVD_CUDA(cudaMemcpyAsync(image, B.b, n * sizeof(float), cudaMemcpyDeviceToHost, B.st));
        } else {
            float* srcp = B.a;
            float* dstp = B.b;
            const float* ks[3] = {cx ? kx : nullptr, cy ? ky : nullptr, cz ? kz : nullptr};
            for (int ax = 0; ax < 3; ++ax) {
                if (!ks[ax]) continue;
                int kd[3] = {1, 1, 1};
                kd[ax] = N;
                convolve_device(dev, srcp, dim, ks[ax], kd, ext, oobv, dstp, B.st);
                std::swap(srcp, dstp);
            }
            MVD_CUDA(cudaStreamSynchronize(B.st));
            MVD_CUDA(cudaMemcpyAsync(image, srcp, n * sizeof(float), cudaMemcpyDeviceToHost, B.st));
        }
        MVD_CUDA(cudaStreamSynchronize(B.st));
    });
    return rc == 0 ? 1 : 0;  // Java boolean
}

#define MVD_SEP_HOOK(N)                                                                                                          \
    extern "C" int convolve_##N(float* image, float* kernelX, float* kernelY, float* kernelZ, int imageW, int imageH, int imageD,   \
                                int convolveX, int convolveY, int convolveZ, int outofbounds, float outofboundsvalue, int devCUDA) { \
        return separable_hook(N, image, kernelX, kernelY, kernelZ, imageW, imageH, imageD, convolveX, convolveY, convolveZ,       \
                              outofbounds, outofboundsvalue, devCUDA);                                                            \
    }
MVD_SEP_HOOK(7)
MVD_SEP_HOOK(15)
MVD_SEP_HOOK(31)
MVD_SEP_HOOK(63)
MVD_SEP_HOOK(127)
#undef MVD_SEP_HOOK

// ---- fused ghost exchange --------------------------------------------------------------------
extern "C" int mvdExportArrays(mvdContext* c, unsigned char psiHandle[64], unsigned char quotientHandle[64]) {
    return guarded([&] {
        require(c && psiHandle && quotientHandle, "mvdExportArrays: null argument");
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
        c->activate();
        cudaIpcMemHandle_t h;
        MVD_CUDA(cudaIpcGetMemHandle(&h, c->psi));
        std::memcpy(psiHandle, &h, 64);
        MVD_CUDA(cudaIpcGetMemHandle(&h, c->quot));
        std::memcpy(quotientHandle, &h, 64);
    });
}

extern "C" int mvdAttachPeer(mvdContext* c, const int peerBrickOffset[3], const int peerBrickDim[3], const void* psi,
                             const void* quotient, int kind, int peerDevice) {
    return guarded([&] {
        require(c && peerBrickOffset && peerBrickDim && psi && quotient, "mvdAttachPeer: null argument");
        require(kind == MVD_PEER_POINTER || kind == MVD_PEER_IPC, "mvdAttachPeer: bad kind");
        require((int)c->peers_psi.size() < kMaxPeers, "mvdAttachPeer: too many neighbours");
        c->activate();
        int pglo[3], pghi[3], padim[3];
        for (int d = 0; d < 3; ++d)
            require(peerBrickDim[d] > 0 && peerBrickOffset[d] >= 0 && peerBrickOffset[d] + peerBrickDim[d] <= c->gdim[d],
                    "mvdAttachPeer: neighbour brick outside the volume");
        ghost_geometry(c->gdim, peerBrickOffset, peerBrickDim, c->ghost, pglo, pghi, padim);
        PeerTarget t;
        std::memset(&t, 0, sizeof(t));
        bool any = true;
        for (int d = 0; d < 3; ++d) {
            // the neighbour wants my voxels inside its ghosted extent: the true ghost width in y and z, the
            // whole (4-aligned) array extent in x so that quads stay whole
            const int want_lo = peerBrickOffset[d] - (d == 0 ? pglo[d] : std::min(c->ghost[d], pglo[d]));
            const int want_hi = peerBrickOffset[d] + peerBrickDim[d] + (d == 0 ? pghi[d] : std::min(c->ghost[d], pghi[d]));
            t.dim[d] = padim[d];
            t.shift[d] = c->boff[d] - (peerBrickOffset[d] - pglo[d]);
            t.lo[d] = std::max(0, want_lo - c->boff[d]);
            t.hi[d] = std::min(c->bdim[d], want_hi - c->boff[d]);
            any = any && t.lo[d] < t.hi[d];
        }
        require(any, "mvdAttachPeer: the neighbour's ghost layers do not reach this brick");
        float* ppsi = nullptr;
        float* pquot = nullptr;
        if (kind == MVD_PEER_IPC) {
            cudaIpcMemHandle_t h;
            void* m = nullptr;
            std::memcpy(&h, psi, 64);
            MVD_CUDA(cudaIpcOpenMemHandle(&m, h, cudaIpcMemLazyEnablePeerAccess));
            c->ipc_mappings.push_back(m);
            ppsi = static_cast<float*>(m);
            std::memcpy(&h, quotient, 64);
            MVD_CUDA(cudaIpcOpenMemHandle(&m, h, cudaIpcMemLazyEnablePeerAccess));
            c->ipc_mappings.push_back(m);
            pquot = static_cast<float*>(m);
        } else {
            if (peerDevice >= 0 && peerDevice != c->device) {
                int can = 0;
                MVD_CUDA(cudaDeviceCanAccessPeer(&can, c->device, peerDevice));
                require(can != 0, "mvdAttachPeer: no peer access between the two devices");
                const cudaError_t e = cudaDeviceEnablePeerAccess(peerDevice, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else MVD_CUDA(e);
            }
            ppsi = const_cast<float*>(static_cast<const float*>(psi));
            pquot = const_cast<float*>(static_cast<const float*>(quotient));
        }
        MVD_CUDA(cudaStreamSynchronize(c->stream));
        t.base = ppsi;
        c->peers_psi.push_back(t);
        t.base = pquot;
        c->peers_quot.push_back(t);
        if (!c->d_peers_psi) {
            MVD_CUDA(cudaMalloc(&c->d_peers_psi, kMaxPeers * sizeof(PeerTarget)));
            MVD_CUDA(cudaMalloc(&c->d_peers_quot, kMaxPeers * sizeof(PeerTarget)));
        }
        MVD_CUDA(cudaMemcpy(c->d_peers_psi, c->peers_psi.data(), c->peers_psi.size() * sizeof(PeerTarget), cudaMemcpyHostToDevice));
        MVD_CUDA(cudaMemcpy(c->d_peers_quot, c->peers_quot.data(), c->peers_quot.size() * sizeof(PeerTarget), cudaMemcpyHostToDevice));
        for (auto& plan : c->plans) partition_for_peers(*plan, c->peers_psi.data(), (int)c->peers_psi.size());
    });
}

// ---- neighbour-only barrier of the fused exchange -----------------------------------------------------
extern "C" int mvdExportFlags(mvdContext* c, unsigned char flagsHandle[64], void** flagsPointer) {
    return guarded([&] {
        require(c, "mvdExportFlags: null context");
        c->activate();
        c->ensure_flags();
        if (flagsHandle) {
            cudaIpcMemHandle_t h;
            MVD_CUDA(cudaIpcGetMemHandle(&h, c->d_flags));
            std::memcpy(flagsHandle, &h, 64);
        }
        if (flagsPointer) *flagsPointer = c->d_flags;
    });
}

extern "C" int mvdAttachPeerFlags(mvdContext* c, const void* peerFlags, int kind, int theirSlotForMe, int mySlotForThem) {
    return guarded([&] {
        require(c && peerFlags, "mvdAttachPeerFlags: null argument");
        require(kind == MVD_PEER_POINTER || kind == MVD_PEER_IPC, "mvdAttachPeerFlags: bad kind");
        require(theirSlotForMe >= 0 && theirSlotForMe < 32 && mySlotForThem >= 0 && mySlotForThem < 32, "mvdAttachPeerFlags: slots are 0 .. 31");
        require(c->peer_flags.n < kMaxPeers, "mvdAttachPeerFlags: too many neighbours");
        c->activate();
        c->ensure_flags();
        unsigned int* p = nullptr;
        if (kind == MVD_PEER_IPC) {
            cudaIpcMemHandle_t h;
            void* m = nullptr;
            std::memcpy(&h, peerFlags, 64);
            MVD_CUDA(cudaIpcOpenMemHandle(&m, h, cudaIpcMemLazyEnablePeerAccess));
            c->ipc_mappings.push_back(m);
            p = static_cast<unsigned int*>(m);
        } else {
            p = const_cast<unsigned int*>(static_cast<const unsigned int*>(peerFlags));
        }
        const int i = c->peer_flags.n++;
        c->peer_flags.theirs[i] = p;
        c->peer_flags.their_slot[i] = theirSlotForMe;
        c->peer_flags.my_slot[i] = mySlotForThem;
    });
}

static int peer_barrier(mvdContext* c, unsigned int value, int phase, const char* what) {
    return guarded([&] {
        require(c, what);
        require(c->peer_flags.n > 0 && c->d_flags, "peer barrier: no neighbour flags attached (mvdAttachPeerFlags)");
        c->activate();
        MVD_CUDA(launch_peer_barrier(c->peer_flags, c->d_flags, value, c->d_flags + 32, phase, c->stream));
        ++g_kernel_launches;
    });
}
extern "C" int mvdPeerSignal(mvdContext* c, unsigned int value) { return peer_barrier(c, value, 0, "mvdPeerSignal: null context"); }
extern "C" int mvdPeerWait(mvdContext* c, unsigned int value) { return peer_barrier(c, value, 1, "mvdPeerWait: null context"); }
extern "C" int mvdPeerBarrier(mvdContext* c, unsigned int value) { return peer_barrier(c, value, 2, "mvdPeerBarrier: null context"); }

// after a synchronisation: did a bounded spin of the peer barrier expire?
static void check_peer_status(mvdContext* c) {
    if (!c->d_flags || c->peer_flags.n == 0) return;
    unsigned int st = 0;
    MVD_CUDA(cudaMemcpyAsync(&st, c->d_flags + 32, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
    MVD_CUDA(cudaStreamSynchronize(c->stream));
    if (st != 0) throw CudaError("peer barrier timed out: a neighbouring brick did not signal within ~10 s (crashed rank, or the ranks' step sequences differ)");
}

extern "C" int mvdDetachPeers(mvdContext* c) {
    return guarded([&] {
        require(c, "mvdDetachPeers: null context");
        c->activate();
        detach_peers(c);
    });
}

extern "C" int mvdMaskNeverUpdated(mvdContext* c) {
    return guarded([&] {
        require(c && !c->views.empty(), "mvdMaskNeverUpdated: no views");
        c->activate();
        wait_all_views(c);
        MVD_CUDA(launch_mask_never_updated(c->view_ptrs(), c->bdim, c->psi, c->adim, c->glo, c->stream));
        ++g_kernel_launches;
        MVD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

// ---- several devices behind one handle, one process ------------------------------------------------
// The reference drives N devices from one JVM with one thread per device (MVDeconFFT.java:426-445).  mvdMulti does the
// same behind the C ABI: the volume is cut into one brick per device (power-of-two grid, longest axis first, x bricks
// float4-aligned), every brick is an mvdContext with ghost layers, the bricks are attached to each other as
// MVD_PEER_POINTER neighbours (peer access over NVLink: the step kernels store ghost voxels into the neighbours), and a
// worker thread per device issues that device's launches.  Between two steps a brick's stream waits for the events its
// neighbours recorded after the previous step (cudaStreamWaitEvent across devices) -- a neighbour-only barrier on the
// devices, no collective, no host synchronisation of the streams; the host threads meet at a barrier per step only so
// that an event is recorded before anybody waits for it.
#include <condition_variable>
#include <functional>
#include <thread>

namespace {
struct HostBarrier {
    std::mutex m;
    std::condition_variable cv;
    int n, waiting = 0;
    long long gen = 0;
    explicit HostBarrier(int n_) : n(n_) {}
    void arrive_and_wait() {
        std::unique_lock<std::mutex> lk(m);
        const long long g = gen;
        if (++waiting == n) { waiting = 0; ++gen; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
};
}  // namespace

struct mvdMulti {
    int n = 0;
    int gdim[3], grid[3], ghost[3];
    std::vector<int> devices;
    std::vector<mvdContext*> bricks;
    std::vector<std::vector<int>> neighbours;
    std::vector<cudaEvent_t> done[2];  // per brick, alternating by step
    long long step = 0;                 // steps issued so far (decides which event of a pair is current)
    bool attached = false;
    int num_views = 0;
    ~mvdMulti() {
        for (size_t i = 0; i < bricks.size(); ++i) {
            if (bricks[i]) {
                cudaSetDevice(bricks[i]->device);
                if (bricks[i]->stream) cudaStreamSynchronize(bricks[i]->stream);
            }
        }
        for (size_t i = 0; i < bricks.size(); ++i) {
            if (!bricks[i]) continue;
            cudaSetDevice(bricks[i]->device);
            for (int k = 0; k < 2; ++k)
                if (i < done[k].size() && done[k][i]) cudaEventDestroy(done[k][i]);
            mvdDestroy(bricks[i]);
        }
    }
    // fn(brick index) on one thread per brick; the first exception (message) is rethrown on the caller's thread
    void parallel(const std::function<void(int)>& fn) {
        std::vector<std::thread> th;
        std::vector<std::string> err(n);
        std::vector<int> kind(n, 0);
        for (int i = 0; i < n; ++i)
            th.emplace_back([&, i] {
                try { fn(i); }
                catch (const CudaError& e) { err[i] = e.what(); kind[i] = 2; cudaGetLastError(); }
                catch (const std::invalid_argument& e) { err[i] = e.what(); kind[i] = 1; }
                catch (const std::exception& e) { err[i] = e.what(); kind[i] = 3; }
            });
        for (auto& t : th) t.join();
        for (int i = 0; i < n; ++i) {
            if (kind[i] == 1) throw std::invalid_argument("brick " + std::to_string(i) + ": " + err[i]);
            if (kind[i]) throw CudaError("brick " + std::to_string(i) + " (device " + std::to_string(devices[i]) + "): " + err[i]);
        }
    }
};

// brick boundaries b_i = floor4(n i / g) (b_0 = 0, b_g = n): x bricks stay float4-aligned (same rule as parallel.brick_of)
static int brick_bound(int i, int g, int n) { return i == 0 ? 0 : (i == g ? n : (int)((long long)n * i / g / 4 * 4)); }

static void multi_attach(mvdMulti* m) {
    if (m->attached || m->n == 1) { m->attached = true; return; }
    for (int b = 0; b < m->n; ++b) {
        mvdContext* c = m->bricks[b];
        for (int r : m->neighbours[b]) {
            mvdContext* p = m->bricks[r];
            if (mvdAttachPeer(c, p->boff, p->bdim, p->psi, p->quot, MVD_PEER_POINTER, p->device) != 0) {
                char buf[512];
                mvdGetLastError(buf, sizeof(buf));
                throw CudaError(std::string("mvdAttachPeer: ") + buf);
            }
        }
    }
    m->attached = true;
}

extern "C" int mvdCreateMulti(mvdMulti** out, const int* devices, int numDevices, const int globalDim[3], const int maxKernelDim[3],
                              double lambda, float minValue) {
    return guarded([&] {
        require(out && devices && globalDim && maxKernelDim, "mvdCreateMulti: null argument");
        *out = nullptr;
        require(numDevices >= 1 && (numDevices & (numDevices - 1)) == 0, "mvdCreateMulti: the number of devices must be a power of two");
        std::unique_ptr<mvdMulti> m(new mvdMulti);
        m->n = numDevices;
        m->devices.assign(devices, devices + numDevices);
        for (int d = 0; d < 3; ++d) { m->gdim[d] = globalDim[d]; m->grid[d] = 1; }
        // the grid whose largest brick needs the least FFT-window volume (bricks that sit badly on the windows cost up to 20 % more
        // work); ties prefer fewer cuts along x (long rows), then along y -- the same rule as parallel.brick_grid
        {
            static const int xy_sizes[] = {32, 64, 128}, z_sizes[] = {16, 32, 64, 128, 256, 512};
            auto axis_cost = [&](int n, int k, int d) {
                const std::vector<AxisTile> t = d < 2 ? plan_axis(n, k, 4, xy_sizes, 3) : plan_axis(n, k, 1, z_sizes, 6);
                if (t.empty()) return 1e300;
                double c = 0;
                for (const auto& a : t) c += a.b + 6;
                return c;
            };
            double best = 1e300;
            for (int gx = 1; gx <= numDevices; gx *= 2)
                for (int gy = 1; gx * gy <= numDevices; gy *= 2) {
                    const int g[3] = {gx, gy, numDevices / (gx * gy)};
                    double cost = 1.0;
                    bool ok = true;
                    for (int d = 0; d < 3 && ok; ++d) {
                        require(maxKernelDim[d] >= 1, "mvdCreateMulti: bad kernel extent");
                        const int ghost = g[d] > 1 ? std::max(maxKernelDim[d] / 2, maxKernelDim[d] - 1 - maxKernelDim[d] / 2) : 0;
                        double worst = 0;
                        for (int i = 0; i < g[d]; ++i) {
                            const int n = brick_bound(i + 1, g[d], m->gdim[d]) - brick_bound(i, g[d], m->gdim[d]);
                            if (n < std::max(1, ghost)) { ok = false; break; }
                            worst = std::max(worst, axis_cost(n, maxKernelDim[d], d));
                        }
                        cost *= worst;
                    }
                    if (ok && cost < best) {  // strict: the first grid in (gx, gy) order wins ties
                        best = cost;
                        for (int d = 0; d < 3; ++d) m->grid[d] = g[d];
                    }
                }
            require(best < 1e299, "mvdCreateMulti: no brick grid fits the kernel halo");
        }
        for (int d = 0; d < 3; ++d) {
            require(maxKernelDim[d] >= 1, "mvdCreateMulti: bad kernel extent");
            m->ghost[d] = m->grid[d] > 1 ? std::max(maxKernelDim[d] / 2, maxKernelDim[d] - 1 - maxKernelDim[d] / 2) : 0;
        }
        m->bricks.assign(numDevices, nullptr);
        m->neighbours.resize(numDevices);
        for (int k = 0; k < 2; ++k) m->done[k].assign(numDevices, nullptr);
        std::vector<std::array<int, 3>> coords(numDevices);
        for (int r = 0; r < numDevices; ++r) {
            const int cx = r % m->grid[0], cy = (r / m->grid[0]) % m->grid[1], cz = r / (m->grid[0] * m->grid[1]);
            coords[r] = {cx, cy, cz};
            int off[3], dim[3];
            for (int d = 0; d < 3; ++d) {
                off[d] = brick_bound(coords[r][d], m->grid[d], m->gdim[d]);
                dim[d] = brick_bound(coords[r][d] + 1, m->grid[d], m->gdim[d]) - off[d];
                require(dim[d] >= std::max(1, m->ghost[d]), "mvdCreateMulti: bricks thinner than the kernel halo");
            }
            if (mvdCreate(&m->bricks[r], devices[r], m->gdim, off, dim, m->ghost, lambda, minValue) != 0) {
                char buf[512];
                mvdGetLastError(buf, sizeof(buf));
                throw CudaError(std::string("mvdCreate: ") + buf);
            }
            MVD_CUDA(cudaSetDevice(devices[r]));
            for (int k = 0; k < 2; ++k) MVD_CUDA(cudaEventCreateWithFlags(&m->done[k][r], cudaEventDisableTiming));
        }
        for (int r = 0; r < numDevices; ++r)
            for (int q = 0; q < numDevices; ++q) {
                if (q == r) continue;
                bool nb = true;
                for (int d = 0; d < 3; ++d) {
                    const int delta = coords[q][d] - coords[r][d];
                    nb = nb && delta >= -1 && delta <= 1 && (delta == 0 || m->ghost[d] > 0);
                }
                if (nb) m->neighbours[r].push_back(q);
            }
        *out = m.release();
    });
}

extern "C" int mvdDestroyMulti(mvdMulti* m) {
    if (!m) return 0;
    return guarded([&] { delete m; });
}

extern "C" int mvdMultiNumBricks(mvdMulti* m) { return m ? m->n : -1; }

extern "C" int mvdMultiBrickInfo(mvdMulti* m, int brick, int* device, int offset[3], int dim[3]) {
    return guarded([&] {
        require(m && brick >= 0 && brick < m->n, "mvdMultiBrickInfo: bad brick");
        if (device) *device = m->devices[brick];
        for (int d = 0; d < 3; ++d) {
            if (offset) offset[d] = m->bricks[brick]->boff[d];
            if (dim) dim[d] = m->bricks[brick]->bdim[d];
        }
    });
}

// img / weight: whole-volume host arrays (x fastest); every device uploads its brick from them
extern "C" int mvdMultiAddView(mvdMulti* m, const float* img, const float* weight, float weightConst, const float* kernel1, const int k1[3],
                               const float* kernel2, const int k2[3]) {
    return guarded([&] {
        require(m && img && kernel1 && kernel2 && k1 && k2, "mvdMultiAddView: null argument");
        for (int d = 0; d < 3; ++d)
            require(m->grid[d] == 1 || (std::max(k1[d], k2[d]) / 2 <= m->ghost[d] && std::max(k1[d], k2[d]) - 1 - std::max(k1[d], k2[d]) / 2 <= m->ghost[d]),
                    "mvdMultiAddView: kernel larger than the maxKernelDim given to mvdCreateMulti");
        m->parallel([&](int b) {
            mvdContext* c = m->bricks[b];
            add_view(c, img, weight, weightConst, kernel1, k1, kernel2, k2, MVD_MEM_HOST, m->gdim, c->boff, nullptr);
        });
        ++m->num_views;
    });
}

extern "C" int mvdMultiSetDirectPath(mvdMulti* m, int enable) {
    return guarded([&] {
        require(m, "mvdMultiSetDirectPath: null handle");
        for (auto* c : m->bricks) c->direct_enabled = enable != 0;
    });
}

// psi = the average of MVDeconvolution.java:105-127 over the WHOLE volume, in every brick (ghost layers included)
extern "C" int mvdMultiInitPsiAverage(mvdMulti* m, double* avgOut) {
    return guarded([&] {
        require(m && m->num_views > 0, "mvdMultiInitPsiAverage: no views");
        std::vector<double> sums(m->n, 0.0);
        std::vector<long long> counts(m->n, 0);
        auto check = [](int rc, const char* what) {
            if (rc != 0) {
                char buf[512];
                mvdGetLastError(buf, sizeof(buf));
                throw CudaError(std::string(what) + ": " + buf);
            }
        };
        m->parallel([&](int b) { check(mvdFirstIterationSums(m->bricks[b], &sums[b], &counts[b]), "mvdFirstIterationSums"); });
        double s = 0.0;
        long long n = 0;
        for (int b = 0; b < m->n; ++b) { s += sums[b]; n += counts[b]; }
        double avg = s / (double)n;
        if (std::isnan(avg)) avg = 0.5;
        if (avgOut) *avgOut = avg;
        m->parallel([&](int b) {
            check(mvdFillPsi(m->bricks[b], (float)avg), "mvdFillPsi");
            check(mvdSynchronize(m->bricks[b]), "mvdSynchronize");
        });
    });
}

// whole-volume host array <-> the bricks' (ghosted) arrays; `ghosts`: also fill the ghost layers from the volume
static void multi_copy_psi(mvdMulti* m, float* vol, bool to_device) {
    m->parallel([&](int b) {
        mvdContext* c = m->bricks[b];
        c->activate();
        int lo[3], n[3];  // volume box covered by the copy, clipped to the volume
        for (int d = 0; d < 3; ++d) {
            const int a0 = to_device ? c->boff[d] - c->glo[d] : c->boff[d];
            const int a1 = to_device ? c->boff[d] + c->bdim[d] + c->ghi[d] : c->boff[d] + c->bdim[d];
            lo[d] = std::max(0, a0);
            n[d] = std::min(c->gdim[d], a1) - lo[d];
        }
        cudaMemcpy3DParms p;
        std::memset(&p, 0, sizeof(p));
        cudaPitchedPtr h = make_cudaPitchedPtr(vol, (size_t)c->gdim[0] * sizeof(float), c->gdim[0], c->gdim[1]);
        cudaPitchedPtr g = make_cudaPitchedPtr(c->psi, (size_t)c->adim[0] * sizeof(float), c->adim[0], c->adim[1]);
        const cudaPos hpos = make_cudaPos((size_t)lo[0] * sizeof(float), lo[1], lo[2]);
        const cudaPos gpos = make_cudaPos((size_t)(lo[0] - (c->boff[0] - c->glo[0])) * sizeof(float), lo[1] - (c->boff[1] - c->glo[1]),
                                          lo[2] - (c->boff[2] - c->glo[2]));
        if (to_device) { p.srcPtr = h; p.srcPos = hpos; p.dstPtr = g; p.dstPos = gpos; p.kind = cudaMemcpyHostToDevice; }
        else { p.srcPtr = g; p.srcPos = gpos; p.dstPtr = h; p.dstPos = hpos; p.kind = cudaMemcpyDeviceToHost; }
        p.extent = make_cudaExtent((size_t)n[0] * sizeof(float), n[1], n[2]);
        MVD_CUDA(cudaMemcpy3DAsync(&p, c->stream));
        MVD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

extern "C" int mvdMultiSetPsi(mvdMulti* m, const float* psi) {
    return guarded([&] {
        require(m && psi, "mvdMultiSetPsi: null argument");
        multi_copy_psi(m, const_cast<float*>(psi), true);  // ghost layers come straight from the caller's volume
    });
}

extern "C" int mvdMultiGetPsi(mvdMulti* m, float* psi) {
    return guarded([&] {
        require(m && psi, "mvdMultiGetPsi: null argument");
        multi_copy_psi(m, psi, false);
    });
}

extern "C" int mvdMultiMaskNeverUpdated(mvdMulti* m) {
    return guarded([&] {
        require(m && m->num_views > 0, "mvdMultiMaskNeverUpdated: no views");
        m->parallel([&](int b) {
            if (mvdMaskNeverUpdated(m->bricks[b]) != 0) {
                char buf[512];
                mvdGetLastError(buf, sizeof(buf));
                throw CudaError(std::string("mvdMaskNeverUpdated: ") + buf);
            }
        });
    });
}

// numIterations passes over all views (MVDeconvolution.runIteration, :333-444) on all devices; sumChange / maxChange:
// numIterations * numViews values each (sum over the bricks / maximum over the bricks), or NULL
extern "C" int mvdMultiRunIterations(mvdMulti* m, int numIterations, double* sumChange, double* maxChange) {
    return guarded([&] {
        require(m && m->num_views > 0, "mvdMultiRunIterations: no views");
        require(numIterations >= 0, "mvdMultiRunIterations: negative iteration count");
        multi_attach(m);
        const int nv = m->num_views, nslots = numIterations * nv;
        HostBarrier barrier(m->n);
        std::atomic<bool> failed{false};
        std::vector<std::vector<double>> sums(m->n, std::vector<double>(std::max(nslots, 1))), maxs(m->n, std::vector<double>(std::max(nslots, 1)));
        const long long step0 = m->step;
        m->parallel([&](int b) {
            mvdContext* c = m->bricks[b];
            std::string err;
            int kind = 0;
            long long step = step0;
            auto one_step = [&](const std::function<void()>& body) {
                // every thread passes every barrier, whatever happens: a failing brick only stops issuing work
                if (!failed.load()) {
                    try {
                        c->activate();
                        if (step > 0)
                            for (int r : m->neighbours[b]) MVD_CUDA(cudaStreamWaitEvent(c->stream, m->done[(step - 1) & 1][r], 0));
                        body();
                        MVD_CUDA(cudaEventRecord(m->done[step & 1][b], c->stream));
                    } catch (const CudaError& e) { err = e.what(); kind = 2; failed.store(true); }
                    catch (const std::exception& e) { err = e.what(); kind = 1; failed.store(true); }
                }
                ++step;
                barrier.arrive_and_wait();  // all events of this step are recorded before anybody waits for them
            };
            try {
                c->activate();
                c->ensure_stats(nslots);
            } catch (const std::exception& e) { err = e.what(); kind = 2; failed.store(true); }
            for (int it = 0; it < numIterations; ++it)
                for (int v = 0; v < nv; ++v) {
                    one_step([&] { step_quotient(c, v); });
                    one_step([&] { step_update(c, v, it * nv + v); });
                }
            if (!failed.load()) {
                try {
                    if (nslots > 0) read_stats(c, 0, nslots, sums[b].data(), maxs[b].data());
                    else MVD_CUDA(cudaStreamSynchronize(c->stream));
                } catch (const std::exception& e) { err = e.what(); kind = 2; failed.store(true); }
            }
            if (kind == 1) throw std::invalid_argument(err);
            if (kind) throw CudaError(err);
        });
        m->step = step0 + 2LL * nslots;
        for (int i = 0; i < nslots; ++i) {
            double s = 0.0, mx = -1.0;
            for (int b = 0; b < m->n; ++b) { s += sums[b][i]; mx = std::max(mx, maxs[b][i]); }
            if (sumChange) sumChange[i] = s;
            if (maxChange) maxChange[i] = mx;
        }
    });
}

// ---- inputs of the loop, produced on the device ----------------------------------------------
namespace {
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    void alloc(size_t bytes) { MVD_CUDA(cudaMalloc(&p, bytes)); }
};
void use_device(int device) {
    int n_dev = 0;
    MVD_CUDA(cudaGetDeviceCount(&n_dev));
    require(device >= 0 && device < n_dev, "no such CUDA device (this library has no CPU path)");
    MVD_CUDA(cudaSetDevice(device));
    Device::get(device);  // checks the architecture, uploads the tables
    static std::mutex m;
    static std::vector<char> done;
    std::lock_guard<std::mutex> lock(m);
    if ((int)done.size() < n_dev) done.resize(n_dev, 0);
    if (!done[device]) {
        MVD_CUDA(upload_blend_table());
        done[device] = 1;
    }
}
}  // namespace

extern "C" int mvdTransformView(int device, const float* src, const int srcDim[3], const double inverse[12], const int outOffset[3],
                                const int outDim[3], const float blendBorder[3], const float blendRange[3], float minValue, float* outImg,
                                float* outWeight, int srcMemKind, int outMemKind) {
    return guarded([&] {
        require(src && srcDim && inverse && outOffset && outDim && outImg, "mvdTransformView: null argument");
        require(!outWeight || (blendBorder && blendRange), "mvdTransformView: weights need blendBorder and blendRange");
        use_device(device);
        TransformArgs a;
        std::memset(&a, 0, sizeof(a));
        for (int i = 0; i < 12; ++i) a.inv[i] = inverse[i];
        for (int d = 0; d < 3; ++d) {
            require(srcDim[d] > 0 && outDim[d] > 0, "mvdTransformView: empty image");
            a.src_dim[d] = srcDim[d];
            a.out_dim[d] = outDim[d];
            a.offset[d] = outOffset[d];
            a.border[d] = blendBorder ? blendBorder[d] : 0.f;
            a.blending[d] = blendRange ? blendRange[d] : 1.f;
        }
        a.min_value = minValue;
        const size_t ns = (size_t)srcDim[0] * srcDim[1] * srcDim[2], no = (size_t)outDim[0] * outDim[1] * outDim[2];
        StreamGuard sg;
        DevBuf dsrc, dimg, dw;
        const float* s = src;
        if (srcMemKind != MVD_MEM_DEVICE) {
            dsrc.alloc(ns * sizeof(float));
            MVD_CUDA(cudaMemcpyAsync(dsrc.p, src, ns * sizeof(float), cudaMemcpyHostToDevice, sg.st));
            s = static_cast<const float*>(dsrc.p);
        }
        float* oi = outImg;
        float* ow = outWeight;
        if (outMemKind != MVD_MEM_DEVICE) {
            dimg.alloc(no * sizeof(float));
            oi = static_cast<float*>(dimg.p);
            if (outWeight) {
                dw.alloc(no * sizeof(float));
                ow = static_cast<float*>(dw.p);
            }
        }
        MVD_CUDA(launch_transform_view(s, a, oi, ow, sg.st));
        ++g_kernel_launches;
        if (outMemKind != MVD_MEM_DEVICE) {
            MVD_CUDA(cudaMemcpyAsync(outImg, oi, no * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
            if (outWeight) MVD_CUDA(cudaMemcpyAsync(outWeight, ow, no * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
        }
        MVD_CUDA(cudaStreamSynchronize(sg.st));
    });
}

extern "C" int mvdNormalizeWeights(int device, float* const* weights, int numViews, long long numVoxels, double osemSpeedup,
                                   int* minOverlappingViews, double* avgOverlappingViews, int memKind) {
    return guarded([&] {
        require(weights && numViews > 0 && numViews <= kMaxViews && numVoxels > 0, "mvdNormalizeWeights: bad argument");
        use_device(device);
        const size_t n = (size_t)numVoxels;
        StreamGuard sg;
        std::vector<DevBuf> bufs(memKind == MVD_MEM_DEVICE ? 0 : numViews);
        WeightPtrs wp;
        wp.n = numViews;
        for (int v = 0; v < numViews; ++v) {
            require(weights[v], "mvdNormalizeWeights: null weight image");
            if (memKind == MVD_MEM_DEVICE) {
                wp.w[v] = weights[v];
            } else {
                bufs[v].alloc(n * sizeof(float));
                MVD_CUDA(cudaMemcpyAsync(bufs[v].p, weights[v], n * sizeof(float), cudaMemcpyHostToDevice, sg.st));
                wp.w[v] = static_cast<float*>(bufs[v].p);
            }
        }
        DevBuf stats;
        stats.alloc(16);
        unsigned long long* d_count = static_cast<unsigned long long*>(stats.p);
        int* d_min = reinterpret_cast<int*>(d_count + 1);
        const unsigned long long zero = 0;
        const int big = numViews;
        MVD_CUDA(cudaMemcpyAsync(d_count, &zero, sizeof(zero), cudaMemcpyHostToDevice, sg.st));
        MVD_CUDA(cudaMemcpyAsync(d_min, &big, sizeof(big), cudaMemcpyHostToDevice, sg.st));
        MVD_CUDA(launch_normalize_weights(wp, n, d_count, d_min, sg.st));
        ++g_kernel_launches;
        if (osemSpeedup != 1.0 && osemSpeedup > 0.0)
            for (int v = 0; v < numViews; ++v) {
                MVD_CUDA(launch_osem(wp.w[v], n, (float)osemSpeedup, sg.st));
                ++g_kernel_launches;
            }
        unsigned long long cnt = 0;
        int mn = 0;
        MVD_CUDA(cudaMemcpyAsync(&cnt, d_count, sizeof(cnt), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaMemcpyAsync(&mn, d_min, sizeof(mn), cudaMemcpyDeviceToHost, sg.st));
        if (memKind != MVD_MEM_DEVICE)
            for (int v = 0; v < numViews; ++v)
                MVD_CUDA(cudaMemcpyAsync(weights[v], wp.w[v], n * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaStreamSynchronize(sg.st));
        if (minOverlappingViews) *minOverlappingViews = mn;
        if (avgOverlappingViews) *avgOverlappingViews = (double)cnt / (double)n;
    });
}

// ---- virtual weights (WeightNormalizer with a sum image + NormalizingRandomAccess) ---------------------
static BlendView make_blend_view(const int srcDim[3], const double inverse[12], const float border[3], const float range[3]) {
    BlendView b;
    for (int i = 0; i < 12; ++i) b.inv[i] = inverse[i];
    for (int d = 0; d < 3; ++d) {
        require(srcDim[d] > 0, "virtual weights: empty view");
        b.dim_minus1[d] = srcDim[d] - 1;
        b.border[d] = border[d];
        b.blending[d] = range[d];
    }
    return b;
}
static VolumeBox make_box(const int outOffset[3], const int outDim[3]) {
    VolumeBox box;
    for (int d = 0; d < 3; ++d) {
        require(outDim[d] > 0, "virtual weights: empty bounding box");
        box.offset[d] = outOffset[d];
        box.dim[d] = outDim[d];
    }
    return box;
}

extern "C" int mvdVirtualWeightSum(int device, int numViews, const int* srcDims, const double* inverses, const float* blendBorders,
                                   const float* blendRanges, const int outOffset[3], const int outDim[3], float* sumWeights,
                                   int* minOverlappingViews, double* avgOverlappingViews, int memKind) {
    return guarded([&] {
        require(numViews > 0 && numViews <= kMaxViews && srcDims && inverses && blendBorders && blendRanges && outOffset && outDim && sumWeights,
                "mvdVirtualWeightSum: bad argument");
        use_device(device);
        const VolumeBox box = make_box(outOffset, outDim);
        const size_t n = (size_t)outDim[0] * outDim[1] * outDim[2];
        std::vector<BlendView> views(numViews);
        for (int v = 0; v < numViews; ++v) views[v] = make_blend_view(srcDims + 3 * v, inverses + 12 * v, blendBorders + 3 * v, blendRanges + 3 * v);
        StreamGuard sg;
        DevBuf dviews, dsum, stats;
        dviews.alloc(views.size() * sizeof(BlendView));
        MVD_CUDA(cudaMemcpyAsync(dviews.p, views.data(), views.size() * sizeof(BlendView), cudaMemcpyHostToDevice, sg.st));
        float* out = sumWeights;
        if (memKind != MVD_MEM_DEVICE) {
            dsum.alloc(n * sizeof(float));
            out = static_cast<float*>(dsum.p);
        }
        stats.alloc(16);
        unsigned long long* d_count = static_cast<unsigned long long*>(stats.p);
        int* d_min = reinterpret_cast<int*>(d_count + 1);
        const unsigned long long zero = 0;
        const int big = numViews;
        MVD_CUDA(cudaMemcpyAsync(d_count, &zero, sizeof(zero), cudaMemcpyHostToDevice, sg.st));
        MVD_CUDA(cudaMemcpyAsync(d_min, &big, sizeof(big), cudaMemcpyHostToDevice, sg.st));
        MVD_CUDA(launch_virtual_weight_sum(static_cast<const BlendView*>(dviews.p), numViews, box, out, d_count, d_min, sg.st));
        ++g_kernel_launches;
        unsigned long long cnt = 0;
        int mn = 0;
        MVD_CUDA(cudaMemcpyAsync(&cnt, d_count, sizeof(cnt), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaMemcpyAsync(&mn, d_min, sizeof(mn), cudaMemcpyDeviceToHost, sg.st));
        if (memKind != MVD_MEM_DEVICE) MVD_CUDA(cudaMemcpyAsync(sumWeights, out, n * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaStreamSynchronize(sg.st));
        if (minOverlappingViews) *minOverlappingViews = mn;
        if (avgOverlappingViews) *avgOverlappingViews = (double)cnt / (double)n;
    });
}

extern "C" int mvdVirtualWeight(int device, const int srcDim[3], const double inverse[12], const float blendBorder[3], const float blendRange[3],
                                const int outOffset[3], const int outDim[3], const float* sumWeights, double osemSpeedup, float* outWeight,
                                int memKind) {
    return guarded([&] {
        require(srcDim && inverse && blendBorder && blendRange && outOffset && outDim && sumWeights && outWeight, "mvdVirtualWeight: null argument");
        use_device(device);
        const VolumeBox box = make_box(outOffset, outDim);
        const BlendView view = make_blend_view(srcDim, inverse, blendBorder, blendRange);
        const size_t n = (size_t)outDim[0] * outDim[1] * outDim[2];
        StreamGuard sg;
        DevBuf dsum, dout;
        const float* sw = sumWeights;
        float* out = outWeight;
        if (memKind != MVD_MEM_DEVICE) {
            dsum.alloc(n * sizeof(float));
            dout.alloc(n * sizeof(float));
            MVD_CUDA(cudaMemcpyAsync(dsum.p, sumWeights, n * sizeof(float), cudaMemcpyHostToDevice, sg.st));
            sw = static_cast<const float*>(dsum.p);
            out = static_cast<float*>(dout.p);
        }
        MVD_CUDA(launch_virtual_weight(view, box, sw, osemSpeedup, out, sg.st));
        ++g_kernel_launches;
        if (memKind != MVD_MEM_DEVICE) MVD_CUDA(cudaMemcpyAsync(outWeight, out, n * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaStreamSynchronize(sg.st));
    });
}

// ---- PSF extraction (ExtractPSF.java) --------------------------------------------------------------
static void invert_affine(const double m[12], double inv[12]) {
    // AffineTransform3D.inverse(): adjugate of the 3x3 part over its determinant, translation -R^-1 t
    const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    require(det != 0.0 && std::isfinite(det), "singular affine transformation");
    const double r[9] = {(e * i - f * h) / det, (c * h - b * i) / det, (b * f - c * e) / det,
                         (f * g - d * i) / det, (a * i - c * g) / det, (c * d - a * f) / det,
                         (d * h - e * g) / det, (b * g - a * h) / det, (a * e - b * d) / det};
    for (int row = 0; row < 3; ++row) {
        for (int col = 0; col < 3; ++col) inv[4 * row + col] = r[3 * row + col];
        inv[4 * row + 3] = -(r[3 * row] * m[3] + r[3 * row + 1] * m[7] + r[3 * row + 2] * m[11]);
    }
}

// ExtractPSF.transformPSF (:304-346): odd size from the bounds of the transformed interval [0, dim-1],
// offset such that the transformed centre voxel (dim / 2) is the centre of the result
static void transformed_psf_geometry(const int dim[3], const double model[12], int new_dim[3], double offset[3]) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int corner = 0; corner < 8; ++corner) {
        const double p[3] = {(corner & 1) ? dim[0] - 1.0 : 0.0, (corner & 2) ? dim[1] - 1.0 : 0.0, (corner & 4) ? dim[2] - 1.0 : 0.0};
        for (int d = 0; d < 3; ++d) {
            const double v = model[4 * d] * p[0] + model[4 * d + 1] * p[1] + model[4 * d + 2] * p[2] + model[4 * d + 3];
            lo[d] = std::min(lo[d], v);
            hi[d] = std::max(hi[d], v);
        }
    }
    const double centre[3] = {(double)(dim[0] / 2), (double)(dim[1] / 2), (double)(dim[2] / 2)};
    for (int d = 0; d < 3; ++d) {
        new_dim[d] = (int)(hi[d] - lo[d]) + 1;
        if (new_dim[d] % 2 == 0) ++new_dim[d];
        const double t = model[4 * d] * centre[0] + model[4 * d + 1] * centre[1] + model[4 * d + 2] * centre[2] + model[4 * d + 3];
        offset[d] = t - (double)(new_dim[d] / 2);
    }
}

extern "C" int mvdExtractPSF(int device, const float* img, const int imgDim[3], const double* locations, int numLocations,
                             const int psfSize[3], int normalize, float* psf, int imgMemKind) {
    return guarded([&] {
        require(img && imgDim && psfSize && psf && (locations || numLocations == 0) && numLocations >= 0, "mvdExtractPSF: bad argument");
        use_device(device);
        PsfArgs a;
        std::memset(&a, 0, sizeof(a));
        for (int d = 0; d < 3; ++d) {
            require(imgDim[d] > 0 && psfSize[d] > 0, "mvdExtractPSF: empty image or PSF");
            a.dim[d] = imgDim[d];
            a.size[d] = psfSize[d];
        }
        const size_t ni = (size_t)imgDim[0] * imgDim[1] * imgDim[2], np = (size_t)psfSize[0] * psfSize[1] * psfSize[2];
        require(np < (1u << 30), "mvdExtractPSF: PSF too large");
        StreamGuard sg;
        DevBuf dimg, dloc, dpsf;
        const float* s = img;
        if (imgMemKind != MVD_MEM_DEVICE) {
            dimg.alloc(ni * sizeof(float));
            MVD_CUDA(cudaMemcpyAsync(dimg.p, img, ni * sizeof(float), cudaMemcpyHostToDevice, sg.st));
            s = static_cast<const float*>(dimg.p);
        }
        dloc.alloc(std::max<size_t>(1, (size_t)numLocations * 3) * sizeof(double));
        if (numLocations) MVD_CUDA(cudaMemcpyAsync(dloc.p, locations, (size_t)numLocations * 3 * sizeof(double), cudaMemcpyHostToDevice, sg.st));
        dpsf.alloc(np * sizeof(float));
        MVD_CUDA(launch_extract_psf(s, a, static_cast<const double*>(dloc.p), numLocations, static_cast<float*>(dpsf.p), sg.st));
        ++g_kernel_launches;
        MVD_CUDA(cudaMemcpyAsync(psf, dpsf.p, np * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaStreamSynchronize(sg.st));
        if (normalize) {  // ExtractPSF.normalize (:277-295): (v - min) / (max - min) in double
            double mn = std::numeric_limits<double>::max(), mx = -std::numeric_limits<double>::max();
            for (size_t i = 0; i < np; ++i) { mn = std::min(mn, (double)psf[i]); mx = std::max(mx, (double)psf[i]); }
            for (size_t i = 0; i < np; ++i) psf[i] = (float)(((double)psf[i] - mn) / (mx - mn));
        }
    });
}

extern "C" int mvdTransformedPSFSize(const int psfDim[3], const double model[12], int newDim[3]) {
    return guarded([&] {
        require(psfDim && model && newDim && psfDim[0] > 0 && psfDim[1] > 0 && psfDim[2] > 0, "mvdTransformedPSFSize: bad argument");
        double offset[3];
        transformed_psf_geometry(psfDim, model, newDim, offset);
    });
}

extern "C" int mvdTransformPSF(int device, const float* psf, const int psfDim[3], const double model[12], const int newDim[3], float* out) {
    return guarded([&] {
        require(psf && psfDim && model && newDim && out, "mvdTransformPSF: null argument");
        use_device(device);
        PsfArgs a;
        std::memset(&a, 0, sizeof(a));
        int want[3];
        transformed_psf_geometry(psfDim, model, want, a.offset);
        for (int d = 0; d < 3; ++d) {
            require(psfDim[d] > 0 && newDim[d] == want[d], "mvdTransformPSF: newDim must come from mvdTransformedPSFSize");
            a.dim[d] = psfDim[d];
            a.size[d] = newDim[d];
        }
        invert_affine(model, a.inv);
        const size_t ns = (size_t)psfDim[0] * psfDim[1] * psfDim[2], no = (size_t)newDim[0] * newDim[1] * newDim[2];
        require(no < (1u << 30), "mvdTransformPSF: result too large");
        StreamGuard sg;
        DevBuf dsrc, dout;
        dsrc.alloc(ns * sizeof(float));
        dout.alloc(no * sizeof(float));
        MVD_CUDA(cudaMemcpyAsync(dsrc.p, psf, ns * sizeof(float), cudaMemcpyHostToDevice, sg.st));
        MVD_CUDA(launch_transform_psf(static_cast<const float*>(dsrc.p), a, static_cast<float*>(dout.p), sg.st));
        ++g_kernel_launches;
        MVD_CUDA(cudaMemcpyAsync(out, dout.p, no * sizeof(float), cudaMemcpyDeviceToHost, sg.st));
        MVD_CUDA(cudaStreamSynchronize(sg.st));
    });
}

// ---- stand-alone convolution ---------------------------------------------------------------
// dims in (x,y,z); src/out host pointers.
static void convolve_device(Device& dev, const float* d_src, const int dim[3], const float* kernel, const int k[3], int ext,
                            float ext_value, float* d_out, cudaStream_t st) {
    auto plan = make_plan(dim, k);
    auto spec = build_spectra(dev, *plan, kernel, st);
    Volume v;
    v.data = d_src;
    for (int d = 0; d < 3; ++d) { v.dim[d] = dim[d]; v.off[d] = 0; v.gmin[d] = 0; v.gdim[d] = dim[d]; }
    v.ext = ext;
    v.ext_value = ext_value;
    EpilogueArgs ep;
    std::memset(&ep, 0, sizeof(ep));
    ep.out = d_out;
    for (int d = 0; d < 3; ++d) { ep.out_dim[d] = dim[d]; ep.out_off[d] = 0; ep.aux_dim[d] = dim[d]; }
    Scratch scratch;
    run_convolution(dev, *plan, *spec, v, EPI_STORE, ep, scratch, st);
    MVD_CUDA(cudaStreamSynchronize(st));  // plan, spectra and scratch die with this frame
}

static void convolve_host(int device, const float* src, const int dim[3], const float* kernel, const int k[3], int ext, float ext_value,
                          float* out) {
    Device& dev = Device::get(device);
    const size_t n = (size_t)dim[0] * dim[1] * dim[2];
    cudaStream_t st = nullptr;
    float *d_src = nullptr, *d_out = nullptr;
    struct Cleanup {
        cudaStream_t& st; float*& a; float*& b;
        ~Cleanup() { if (a) cudaFree(a); if (b) cudaFree(b); if (st) cudaStreamDestroy(st); }
    } cleanup{st, d_src, d_out};
    MVD_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    MVD_CUDA(cudaMalloc(&d_src, n * sizeof(float)));
    MVD_CUDA(cudaMalloc(&d_out, n * sizeof(float)));
    MVD_CUDA(cudaMemcpyAsync(d_src, src, n * sizeof(float), cudaMemcpyHostToDevice, st));
    convolve_device(dev, d_src, dim, kernel, k, ext, ext_value, d_out, st);
    MVD_CUDA(cudaMemcpyAsync(out, d_out, n * sizeof(float), cudaMemcpyDeviceToHost, st));
    MVD_CUDA(cudaStreamSynchronize(st));
}

extern "C" int mvdConvolveDevice(int device, const float* dSrc, const int dim[3], const float* kernel, const int kernelDim[3],
                                 int extMode, float extValue, float* dOut, void* cudaStream) {
    return guarded([&] {
        require(dSrc && dim && kernel && kernelDim && dOut, "mvdConvolveDevice: null argument");
        require(extMode >= 0 && extMode <= 2, "mvdConvolveDevice: bad extension mode");
        require(dSrc != dOut, "mvdConvolveDevice: in-place convolution is not possible (tiles overlap)");
        Device& dev = Device::get(device);
        convolve_device(dev, dSrc, dim, kernel, kernelDim, extMode, extValue, dOut, (cudaStream_t)cudaStream);
    });
}

extern "C" int mvdConvolve(int device, const float* src, const int dim[3], const float* kernel, const int kernelDim[3], int extMode,
                           float extValue, float* out) {
    return guarded([&] {
        require(src && dim && kernel && kernelDim && out, "mvdConvolve: null argument");
        require(extMode >= 0 && extMode <= 2, "mvdConvolve: bad extension mode");
        convolve_host(device, src, dim, kernel, kernelDim, extMode, extValue, out);
    });
}

extern "C" int mvdPlanTiles(const int dim[3], const int kernelDim[3], int* out, int maxTiles) {
    int count = -1;
    guarded([&] {
        require(dim && kernelDim, "mvdPlanTiles: null argument");
        static const int xy[] = {32, 64, 128};
        static const int zs[] = {16, 32, 64, 128, 256, 512};
        std::vector<AxisTile> ax[3] = {plan_axis(dim[0], kernelDim[0], 4, xy, 3), plan_axis(dim[1], kernelDim[1], 4, xy, 3),
                                       plan_axis(dim[2], kernelDim[2], 1, zs, 6)};
        require(!ax[0].empty() && !ax[1].empty() && !ax[2].empty(), "mvdPlanTiles: kernel too large");
        int n = 0;
        for (const auto& tz : ax[2])
            for (const auto& ty : ax[1])
                for (const auto& tx : ax[0]) {
                    if (out && n < maxTiles) {
                        const AxisTile* t[3] = {&tx, &ty, &tz};
                        for (int d = 0; d < 3; ++d) {
                            out[n * 12 + d] = t[d]->b;
                            out[n * 12 + 3 + d] = t[d]->o;
                            out[n * 12 + 6 + d] = t[d]->e;
                            out[n * 12 + 9 + d] = t[d]->n;
                        }
                    }
                    ++n;
                }
        count = n;
    });
    return count;
}

// ---- reference block tiling, bit-exact (BlockGeneratorFixedSizePrecise.java:25-101) ----------
extern "C" long long mvdDivideIntoBlocks(const long long* imgSize, const long long* kernelSize, const long long* blockSize, int nd,
                                         long long* out, long long maxBlocks) {
    if (!imgSize || !kernelSize || !blockSize || nd <= 0 || nd > 8) return -1;
    long long eff[8], loc[8], num[8], total = 1;
    for (int d = 0; d < nd; ++d) {
        eff[d] = blockSize[d] - kernelSize[d] + 1;  // :36
        if (eff[d] <= 0) return 0;                  // :38-42 (returns null)
        loc[d] = kernelSize[d] / 2;                 // :44
        num[d] = imgSize[d] / eff[d];               // :52
        if (imgSize[d] % eff[d] != 0) ++num[d];     // :55-56
        total *= num[d];
    }
    if (!out) return total;
    if (total > maxBlocks) return -1;
    long long cur[8] = {0};
    for (long long b = 0; b < total; ++b) {  // dimension 0 fastest (LocalizingZeroMinIntervalIterator, :68-76)
        long long* o = out + b * 5 * nd;
        for (int d = 0; d < nd; ++d) {
            const long long eoff = cur[d] * eff[d];  // :85
            long long esize = eff[d];
            if (eoff + esize > imgSize[d]) esize = imgSize[d] - eoff;  // :88-89
            o[0 * nd + d] = blockSize[d];
            o[1 * nd + d] = eoff - kernelSize[d] / 2;  // :86
            o[2 * nd + d] = esize;
            o[3 * nd + d] = eoff;
            o[4 * nd + d] = loc[d];
        }
        for (int d = 0; d < nd; ++d) {
            if (++cur[d] < num[d]) break;
            cur[d] = 0;
        }
    }
    return total;
}

// ---- legacy JNA symbols ----------------------------------------------------------------------
extern "C" int getNumDevicesCUDA(void) {
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e == cudaErrorNoDevice) return 0;
    if (e != cudaSuccess) { cudaGetLastError(); return -1; }  // CUDATools.java:39-43 "driver crashed"
    return n;
}

static bool device_prop(int dev, cudaDeviceProp& p) {
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) { cudaGetLastError(); return false; }
    return true;
}

extern "C" int getCUDAcomputeCapabilityMajorVersion(int dev) {
    cudaDeviceProp p;
    return device_prop(dev, p) ? p.major : -1;
}
extern "C" int getCUDAcomputeCapabilityMinorVersion(int dev) {
    cudaDeviceProp p;
    return device_prop(dev, p) ? p.minor : -1;
}
extern "C" void getNameDeviceCUDA(int dev, char* name) {
    if (!name) return;
    cudaDeviceProp p;
    if (!device_prop(dev, p)) { name[0] = 0; return; }
    std::strncpy(name, p.name, 255);  // caller's buffer is 256 bytes (CUDATools.java:55)
    name[255] = 0;
}
extern "C" long long getMemDeviceCUDA(int dev) {
    cudaDeviceProp p;
    return device_prop(dev, p) ? (long long)p.totalGlobalMem : -1;
}
extern "C" long long getFreeMemDeviceCUDA(int dev) {
    size_t f = 0, t = 0;
    if (cudaSetDevice(dev) != cudaSuccess || cudaMemGetInfo(&f, &t) != cudaSuccess) { cudaGetLastError(); return -1; }
    return (long long)f;
}

// State the drop-in symbol keeps per device between calls (guarded by Device::legacy_mutex): the Java side calls
// it once per block, convolution, view and iteration with the same few kernels (MVDeconFFTThreads.java:52-94), so
// the block buffers, the tile plan and the kernel spectra are built once and reused.  Never destroyed: the CUDA
// context may be gone by the time static destructors run.
namespace {
struct LegacyEntry {
    int dim[3], k[3];
    std::vector<float> kernel;  // compared byte-wise: the caller owns its array and may change it
    std::unique_ptr<ConvPlan> plan;
    std::unique_ptr<KernelSpectra> spec;
    unsigned long long stamp = 0;
};
struct LegacyState {
    cudaStream_t st = nullptr;
    float* d_src = nullptr;
    float* d_out = nullptr;
    size_t cap = 0;
    Scratch scratch;
    std::vector<LegacyEntry> cache;
    unsigned long long clock = 0;
    static constexpr size_t kMaxEntries = 32;
};
std::mutex g_legacy_map_mutex;
std::map<int, LegacyState*> g_legacy;

LegacyState& legacy_state(int dev) {
    std::lock_guard<std::mutex> lk(g_legacy_map_mutex);
    LegacyState*& s = g_legacy[dev];
    if (!s) s = new LegacyState();
    return *s;
}
}  // namespace

static int legacy_convolve(const float* im, const int* imDim, const float* kernel, const int* kernelDim, int dev, float* out) {
    return guarded([&] {
        require(im && imDim && kernel && kernelDim && out, "convolution3DfftCUDA: null argument");
        // the reference passes (z,y,x) (MVDeconFFTThreads.java:136-144)
        const int dim[3] = {imDim[2], imDim[1], imDim[0]};
        const int k[3] = {kernelDim[2], kernelDim[1], kernelDim[0]};
        for (int d = 0; d < 3; ++d) require(dim[d] > 0 && k[d] > 0, "convolution3DfftCUDA: non-positive size");
        Device& device = Device::get(dev);
        std::lock_guard<std::mutex> lk(device.legacy_mutex);
        LegacyState& s = legacy_state(dev);
        const size_t n = (size_t)dim[0] * dim[1] * dim[2], kn = (size_t)k[0] * k[1] * k[2];
        if (!s.st) MVD_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
        if (s.cap < n) {
            if (s.d_src) cudaFree(s.d_src);
            if (s.d_out) cudaFree(s.d_out);
            s.d_src = s.d_out = nullptr;
            s.cap = 0;
            MVD_CUDA(cudaMalloc(&s.d_src, n * sizeof(float)));
            MVD_CUDA(cudaMalloc(&s.d_out, n * sizeof(float)));
            s.cap = n;
        }
        // the block goes up while the kernel is looked up (or its spectra are built on a stream of their own)
        MVD_CUDA(cudaMemcpyAsync(s.d_src, im, n * sizeof(float), cudaMemcpyHostToDevice, s.st));
        LegacyEntry* e = nullptr;
        for (auto& c : s.cache)
            if (std::memcmp(c.dim, dim, sizeof(dim)) == 0 && std::memcmp(c.k, k, sizeof(k)) == 0 &&
                std::memcmp(c.kernel.data(), kernel, kn * sizeof(float)) == 0) {
                e = &c;
                break;
            }
        if (!e) {
            if (s.cache.size() >= LegacyState::kMaxEntries) {
                // the spectra of the least recently used kernel may still be read by nothing: every call synchronises
                auto lru = std::min_element(s.cache.begin(), s.cache.end(), [](const LegacyEntry& a, const LegacyEntry& b) { return a.stamp < b.stamp; });
                s.cache.erase(lru);
            }
            LegacyEntry ne;
            std::memcpy(ne.dim, dim, sizeof(dim));
            std::memcpy(ne.k, k, sizeof(k));
            ne.kernel.assign(kernel, kernel + kn);
            ne.plan = make_plan(dim, k);
            StreamGuard aux;
            ne.spec = build_spectra(device, *ne.plan, kernel, aux.st);
            s.cache.push_back(std::move(ne));
            e = &s.cache.back();
        }
        e->stamp = ++s.clock;
        Volume v;
        v.data = s.d_src;
        for (int d = 0; d < 3; ++d) { v.dim[d] = dim[d]; v.off[d] = 0; v.gmin[d] = 0; v.gdim[d] = dim[d]; }
        v.ext = EXT_ZERO;
        v.ext_value = 0.f;
        EpilogueArgs ep;
        std::memset(&ep, 0, sizeof(ep));
        ep.out = s.d_out;
        for (int d = 0; d < 3; ++d) { ep.out_dim[d] = dim[d]; ep.out_off[d] = 0; ep.aux_dim[d] = dim[d]; }
        run_convolution(device, *e->plan, *e->spec, v, EPI_STORE, ep, s.scratch, s.st);
        // `out` may be the caller's block: nothing is written to it before the device has finished without error
        MVD_CUDA(cudaStreamSynchronize(s.st));
        MVD_CUDA(cudaMemcpyAsync(out, s.d_out, n * sizeof(float), cudaMemcpyDeviceToHost, s.st));
        MVD_CUDA(cudaStreamSynchronize(s.st));
    });
}

extern "C" void convolution3DfftCUDAInPlace(float* im, int* imDim, float* kernel, int* kernelDim, int devCUDA) {
    if (!im || !imDim) { latch(1, "convolution3DfftCUDAInPlace: null argument"); return; }
    // on failure im is untouched (the result is copied back only after the device finished without error)
    legacy_convolve(im, imDim, kernel, kernelDim, devCUDA, im);
}

extern "C" float* convolution3DfftCUDA(float* im, int* imDim, float* kernel, int* kernelDim, int devCUDA) {
    if (!im || !imDim) { latch(1, "convolution3DfftCUDA: null argument"); return nullptr; }
    const size_t n = (size_t)imDim[0] * imDim[1] * imDim[2];
    float* out = (float*)std::malloc(n * sizeof(float));
    if (!out) { latch(3, "out of host memory"); return nullptr; }
    if (legacy_convolve(im, imDim, kernel, kernelDim, devCUDA, out) != 0) { std::free(out); return nullptr; }
    return out;
}

// ---- allocation-free separable convolution ----------------------------------------------------------
// No counterpart in the reference; added for callers that cannot allocate
// device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernels as
// convolve_7 / 15 / 31 (sep_xy_kernel -> sep_z_kernel, sep_kernels.cu); the intermediate of the fused x / y pass lives
// in the file-scope array below instead of a cudaMalloc'ed buffer.  Every GPU that loads the library holds the array
// (MVD_NOALLOC_MAX_VOXELS floats = 32 MiB) until the process exits.
//
// In host code `g_noalloc_scratch` is a host-side stand-in, NOT a device address: the address the kernels get comes
// from cudaGetSymbolAddress, resolved once per device with that device current (noalloc_scratch_ptr).
// Every call shares the array and it keeps the previous call's contents.  Nothing in it has to start at zero: the x / y
// pass writes every element the z pass reads (all dim[0] * dim[1] * dim[2] of them) before the z pass starts, in stream
// order.  Calls that overlap on the GPU (different streams, or different host threads on the same device) RACE on it.
static __device__ __align__(256) float g_noalloc_scratch[MVD_NOALLOC_MAX_VOXELS];

static float* noalloc_scratch_ptr(int device) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static float* ptrs[kMaxDev] = {};
    require(device >= 0 && device < kMaxDev, "mvdSeparableConvolveNoAlloc: device index out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[device]) {
        int cur = -1;
        MVD_CUDA(cudaGetDevice(&cur));
        require(cur == device, "mvdSeparableConvolveNoAlloc: internal error, the device is not current");
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            throw CudaError(std::string("cudaGetSymbolAddress(g_noalloc_scratch) on device ") + std::to_string(device) + ": " + cudaGetErrorString(e));
        ptrs[device] = static_cast<float*>(p);
    }
    return ptrs[device];
}

extern "C" int mvdSeparableConvolveNoAlloc(int device, const float* dSrc, float* dOut, const int dim[3], const float* kernelX,
                                           const float* kernelY, const float* kernelZ, int kernelLen, int outofbounds,
                                           float outofboundsvalue, void* cudaStream) {
    return guarded([&] {
        require(dSrc && dOut && dim, "mvdSeparableConvolveNoAlloc: null argument");
        require(dSrc != dOut, "mvdSeparableConvolveNoAlloc: dSrc and dOut must be different arrays");
        require(dim[0] > 0 && dim[1] > 0 && dim[2] > 0, "mvdSeparableConvolveNoAlloc: bad image size");
        require(kernelLen >= 1 && kernelLen <= kSepMaxTaps, "mvdSeparableConvolveNoAlloc: kernelLen must be 1 .. 31 (direct kernels only)");
        require(outofbounds >= 0 && outofbounds <= 2, "mvdSeparableConvolveNoAlloc: outofbounds must be 0 (zero), 1 (value) or 2 (extend last pixel)");
        const unsigned long long vox = (unsigned long long)dim[0] * (unsigned long long)dim[1] * (unsigned long long)dim[2];
        if (vox > (unsigned long long)MVD_NOALLOC_MAX_VOXELS)
            throw std::invalid_argument("mvdSeparableConvolveNoAlloc: " + std::to_string(vox) + " voxels exceed the maximum of " +
                                        std::to_string((unsigned long long)MVD_NOALLOC_MAX_VOXELS) +
                                        " (MVD_NOALLOC_MAX_VOXELS, the size of the library's static scratch array); use convolve_N or mvdConvolveDevice");
        int n_dev = 0;
        MVD_CUDA(cudaGetDeviceCount(&n_dev));
        require(device >= 0 && device < n_dev, "mvdSeparableConvolveNoAlloc: no such CUDA device (this library has no CPU path)");
        MVD_CUDA(cudaSetDevice(device));
        float* tmp = noalloc_scratch_ptr(device);
        Volume v;
        v.data = dSrc;
        for (int i = 0; i < 3; ++i) { v.dim[i] = dim[i]; v.off[i] = 0; v.gmin[i] = 0; v.gdim[i] = dim[i]; }
        v.ext = outofbounds == 0 ? EXT_ZERO : (outofbounds == 1 ? EXT_CONST : EXT_CLAMP);
        v.ext_value = outofboundsvalue;
        const float one = 1.f;
        SepTaps taps;
        std::memset(&taps, 0, sizeof(taps));
        const int kmax = sep_kmax_for(kernelLen, kernelLen, kernelLen);
        make_window_taps(kernelX ? kernelX : &one, kernelX ? kernelLen : 1, kmax, taps.x);
        make_window_taps(kernelY ? kernelY : &one, kernelY ? kernelLen : 1, kmax, taps.y);
        make_window_taps(kernelZ ? kernelZ : &one, kernelZ ? kernelLen : 1, kmax, taps.z);
        taps.per_pass_fill = 1;  // three 1-D passes, each with the plain outside value, like convolve_N
        EpilogueArgs ep;
        std::memset(&ep, 0, sizeof(ep));
        ep.out = dOut;
        for (int i = 0; i < 3; ++i) { ep.out_dim[i] = dim[i]; ep.aux_dim[i] = dim[i]; }
        // two launches on the caller's stream, nothing else: no allocation, no copy, no synchronisation
        MVD_CUDA(launch_separable(v, dim, taps, kmax, EPI_STORE, ep, tmp, static_cast<cudaStream_t>(cudaStream)));
        g_kernel_launches += 2;
    });
}

// ---- kernel preparation behind the C ABI ------------------------------------------
// MVDeconInput.init / MVDeconFFT.init (MVDeconInput.java:43-49, MVDeconFFT.java:162-333) for callers without the Java
// (or Python) host: kernel1 = normImg(psf), kernel2 by PSFTYPE.  Host arithmetic as in the reference, statement by
// statement; the small zero-extended convolutions of the compound kernels run through the device engine.
namespace {
// correctly rounded sum of doubles (Shewchuk's exact partials; what the oracle's math.fsum computes).  The reference
// accumulates in mpicbg.util.RealSum, a cascaded compensated sum that is exact for these sizes as well.
double exact_sum(const std::vector<double>& x) {
    std::vector<double> partials;
    for (double v : x) {
        size_t i = 0;
        for (double y : partials) {
            if (std::fabs(v) < std::fabs(y)) std::swap(v, y);
            const double hi = v + y;
            const double lo = y - (hi - v);
            if (lo != 0.0) partials[i++] = lo;
            v = hi;
        }
        partials.resize(i);
        partials.push_back(v);
    }
    // sum the partials from the top, correcting the final rounding (as CPython's math.fsum does)
    double hi = 0.0;
    int n = (int)partials.size();
    if (n > 0) {
        hi = partials[--n];
        double lo = 0.0;
        while (n > 0) {
            const double x1 = hi, y = partials[--n];
            hi = x1 + y;
            const double yr = hi - x1;
            lo = y - yr;
            if (lo != 0.0) break;
        }
        if (n > 0 && ((lo < 0.0 && partials[n - 1] < 0.0) || (lo > 0.0 && partials[n - 1] > 0.0))) {
            const double y = lo * 2.0, x1 = hi + y;
            if (y == x1 - hi) hi = x1;
        }
    }
    return hi;
}
// AdjustInput.sumImg (AdjustInput.java:44-100): 2 * numThreads portions (FusionHelper.divideIntoPortions,
// FusionHelper.java:236-260); the partial sum of the first portion is counted twice (:93-97) -- kept.
double sum_img_quirk(const std::vector<float>& img, int num_threads) {
    const size_t n = img.size(), parts = (size_t)num_threads * 2, chunk = n / parts, mod = n % parts;
    std::vector<double> sums;
    for (size_t i = 0; i < parts; ++i) {
        const size_t start = i * chunk, len = chunk + (i == parts - 1 ? mod : 0);
        std::vector<double> v(len);
        for (size_t j = 0; j < len; ++j) v[j] = (double)img[start + j];
        sums.push_back(exact_sum(v));
    }
    std::vector<double> all;
    all.push_back(sums[0]);
    all.insert(all.end(), sums.begin(), sums.end());
    return exact_sum(all);
}
void norm_img(std::vector<float>& img, int num_threads) {  // AdjustInput.java:29-36: t = (float)((double) t / sum)
    const double s = sum_img_quirk(img, num_threads);
    for (float& t : img) t = (float)((double)t / s);
}
// Mirror.mirror along every dimension (Mirror.java:31-108, MVDeconFFT.java:315-323); for an even extent the two middle
// positions are swapped twice, i.e. stay where they are (Mirror.java:72)
std::vector<float> inverted_kernel(const std::vector<float>& k, const int d[3]) {
    std::vector<float> cur = k, nxt(k.size());
    for (int ax = 0; ax < 3; ++ax) {
        const int n = d[ax];
        std::vector<int> idx(n);
        for (int i = 0; i < n; ++i) idx[i] = n - 1 - i;
        if (n % 2 == 0) { idx[n / 2 - 1] = n / 2 - 1; idx[n / 2] = n / 2; }
        for (int z = 0; z < d[2]; ++z)
            for (int y = 0; y < d[1]; ++y)
                for (int x = 0; x < d[0]; ++x) {
                    const int c[3] = {x, y, z};
                    int s[3] = {x, y, z};
                    s[ax] = idx[c[ax]];
                    nxt[((size_t)z * d[1] + y) * d[0] + x] = cur[((size_t)s[2] * d[1] + s[1]) * d[0] + s[0]];
                }
        cur.swap(nxt);
    }
    return cur;
}
}  // namespace

static void convolve_host(int device, const float* src, const int dim[3], const float* kernel, const int k[3], int ext, float ext_value,
                          float* out);

extern "C" int mvdPrepareKernels(int device, int numViews, const float* const* psfs, const int* psfDims, int psfType, int numThreads,
                                 float* const* kernel1, float* const* kernel2) {
    return guarded([&] {
        require(numViews >= 1 && psfs && psfDims && kernel1 && kernel2, "mvdPrepareKernels: null argument");
        require(psfType >= 0 && psfType <= 3, "mvdPrepareKernels: psfType must be 0 OPTIMIZATION_II, 1 OPTIMIZATION_I, 2 EFFICIENT_BAYESIAN, 3 INDEPENDENT");
        require(numThreads >= 1, "mvdPrepareKernels: numThreads must be positive");
        const int V = numViews;
        std::vector<std::vector<float>> k1(V);
        for (int v = 0; v < V; ++v) {
            const int* d = psfDims + 3 * v;
            require(psfs[v] && kernel1[v] && kernel2[v] && d[0] > 0 && d[1] > 0 && d[2] > 0, "mvdPrepareKernels: bad view");
            k1[v].assign(psfs[v], psfs[v] + (size_t)d[0] * d[1] * d[2]);
        }
        auto conv0 = [&](const std::vector<float>& in, const int* din, const std::vector<float>& k, const int* dk) {
            // FFTConvolution(Views.extendZero(input), input, Views.extendZero(kernel), kernel, output): same extent as the input
            std::vector<float> out(in.size());
            convolve_host(device, in.data(), din, k.data(), dk, EXT_ZERO, 0.f, out.data());
            return out;
        };
        // views are initialised in order: the kernels of later views are still un-normalised when an earlier view combines
        // them (MVDeconInput.java:43-49)
        for (int v = 0; v < V; ++v) {
            const int* dv = psfDims + 3 * v;
            norm_img(k1[v], numThreads);                                           // MVDeconFFT.java:165
            std::vector<float> k2;
            if (V == 1 || psfType == 3) {                                          // :174-180
                k2 = inverted_kernel(k1[v], dv);
            } else if (psfType == 2) {                                             // EFFICIENT_BAYESIAN :181-244
                std::vector<float> tmp = inverted_kernel(k1[v], dv);
                for (int w = 0; w < V; ++w) {
                    if (w == v) continue;
                    const int* dw = psfDims + 3 * w;
                    std::vector<float> out = conv0(inverted_kernel(k1[v], dv), dv, k1[w], dw);
                    out = conv0(out, dv, inverted_kernel(k1[w], dw), dw);
                    for (size_t i = 0; i < tmp.size(); ++i) tmp[i] = out[i] * tmp[i];
                }
                norm_img(tmp, numThreads);
                k2 = tmp;
            } else if (psfType == 1) {                                             // OPTIMIZATION_I :245-291
                std::vector<float> tmp = k1[v];
                for (int w = 0; w < V; ++w) {
                    if (w == v) continue;
                    const int* dw = psfDims + 3 * w;
                    const std::vector<float> out = conv0(k1[v], dv, inverted_kernel(k1[w], dw), dw);
                    for (size_t i = 0; i < tmp.size(); ++i) tmp[i] = out[i] * tmp[i];
                }
                norm_img(tmp, numThreads);
                k2 = inverted_kernel(tmp, dv);
            } else {                                                               // OPTIMIZATION_II :292-302
                std::vector<float> e = k1[v];
                for (int i = 1; i < V; ++i)                                        // computeExponentialKernel :305-333
                    for (size_t j = 0; j < e.size(); ++j) e[j] = e[j] * k1[v][j];
                norm_img(e, numThreads);
                k2 = inverted_kernel(e, dv);
            }
            std::memcpy(kernel1[v], k1[v].data(), k1[v].size() * sizeof(float));
            std::memcpy(kernel2[v], k2.data(), k2.size() * sizeof(float));
        }
    });
}

// CUDASeparableConvolution.convolutionCPU (CUDASeparableConvolution.java:19-21; declared by the reference, never called):
// in-place separable convolution with kernels of RADIUS kernelR* (2 R + 1 taps; a null kernel leaves its axis alone).  This
// library has no CPU code: the same result is computed on device 0 by the kernels behind convolve_N.
static int separable_hook(int N, float* image, const float* kx, const float* ky, const float* kz, int w, int h, int d, int cx, int cy,
                          int cz, int oob, float oobv, int dev_id);
extern "C" void convolutionCPU(float* image, float* kernelX, float* kernelY, float* kernelZ, int kernelRX, int kernelRY, int kernelRZ,
                               int imageW, int imageH, int imageD, int outofbounds, float outofboundsvalue) {
    if (kernelRX < 0 || kernelRY < 0 || kernelRZ < 0) { latch(1, "convolutionCPU: negative kernel radius"); return; }
    const int r = std::max(kernelRX, std::max(kernelRY, kernelRZ)), N = 2 * r + 1;
    auto centred = [&](const float* k, int rk) {  // kernels shorter than N are centred, like getFloatKernelPadded does for convolve_N
        std::vector<float> p(N, 0.f);
        if (k) std::memcpy(p.data() + (r - rk), k, (size_t)(2 * rk + 1) * sizeof(float));
        return p;
    };
    const std::vector<float> px = centred(kernelX, kernelRX), py = centred(kernelY, kernelRY), pz = centred(kernelZ, kernelRZ);
    separable_hook(N, image, px.data(), py.data(), pz.data(), imageW, imageH, imageD, kernelX != nullptr, kernelY != nullptr, kernelZ != nullptr,
                   outofbounds, outofboundsvalue, 0);
}
