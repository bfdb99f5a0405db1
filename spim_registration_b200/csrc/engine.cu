// Host side of the tiled convolution engine (see engine.h).
#include "engine.h"
#include "volume_kernels.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>

namespace mvd {

std::atomic<long long> g_kernel_launches{0};
static int env_flag(const char* name, int dflt) {
    const char* s = std::getenv(name);
    return s ? std::atoi(s) : dflt;
}
std::atomic<int> g_use_z_tma{env_flag("MVD_Z_TMA", 0)};
std::atomic<int> g_use_inv_ring{env_flag("MVD_INV_RING", 2)};
// 0: register-loading z pass; 1: warp-local lines for 512-point windows (z_warp_kernel); 2: also for 256-point windows;
// 3: lines on part of a warp, two passes (z_part_kernel) for 512-point windows; 4: also for 256-point windows
std::atomic<int> g_use_z_warp{env_flag("MVD_Z_WARP", 4)};
static bool z_warp_covers(int bz) {
    const int g = g_use_z_warp.load();
    return g != 0 && (bz == 512 || g == 2 || g == 4);
}

void cuda_check(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw CudaError(std::string(what) + ": " + cudaGetErrorString(e));
}

// ---- planner ---------------------------------------------------------------------------
static inline int round_up(int v, int a) { return (v + a - 1) / a * a; }

std::vector<AxisTile> plan_axis(int n, int k, int align, const int* sizes, int num_sizes) {
    const int hl = k - 1 - k / 2, hr = k / 2;
    const int lo = round_up(hl, align);
    std::vector<int> eff(num_sizes);
    bool any = false;
    for (int i = 0; i < num_sizes; ++i) {
        int e = sizes[i] - lo - hr;
        e = e / align * align;
        eff[i] = e;
        any |= e > 0;
    }
    if (!any || n <= 0) return {};
    // dynamic programme over the first uncovered voxel; cost = sum of window lengths (+ small per-window charge)
    const long long per_tile = 6;
    std::vector<long long> cost(n + 1, 0);
    std::vector<int> pick(n + 1, -1);
    for (int pos = n - 1; pos >= 0; --pos) {
        long long best = -1;
        int bi = -1;
        for (int i = num_sizes - 1; i >= 0; --i) {  // ties prefer the larger window
            if (eff[i] <= 0) continue;
            const long long c = sizes[i] + per_tile + cost[std::min(n, pos + eff[i])];
            if (best < 0 || c < best) { best = c; bi = i; }
        }
        cost[pos] = best;
        pick[pos] = bi;
    }
    std::vector<int> chosen;
    for (int pos = 0; pos < n;) {
        chosen.push_back(pick[pos]);
        pos += eff[pick[pos]];
    }
    std::sort(chosen.begin(), chosen.end(), [&](int a, int b) { return sizes[a] > sizes[b]; });
    std::vector<AxisTile> out;
    int e = 0;
    for (int i : chosen) {
        if (e >= n) break;
        AxisTile t;
        t.b = sizes[i];
        t.e = e;
        t.o = e - lo;
        t.n = std::min(eff[i], n - e);
        out.push_back(t);
        e += eff[i];
    }
    return out;
}

ConvPlan::~ConvPlan() {
    for (auto& c : classes)
        if (c.d_tiles) cudaFree(c.d_tiles);
}

bool same_plan(const ConvPlan& p, const int dim[3], const int k[3]) {
    for (int d = 0; d < 3; ++d)
        if (p.dim[d] != dim[d] || p.k[d] != k[d]) return false;
    return true;
}

static const int kXYSizes[] = {32, 64, 128, 256};
static const int kZSizes[] = {16, 32, 64, 128, 256, 512};

// x, y: windows up to 128 points (two 256-thread CTAs per SM: measured faster per output voxel than 256-point
// windows at one 512-thread CTA per SM in spite of their smaller halo share, 31.35 vs 31.43 ms per iteration at
// K = 31); 256-point windows only for kernels that do not fit 128, or on request (MVD_XMAX / MVD_YMAX = 256).
static std::vector<AxisTile> plan_axes(const int dim[3], const int k[3], int d) {
    if (d < 2) {
        int n = 3;
        if (const char* s = std::getenv(d == 0 ? "MVD_XMAX" : "MVD_YMAX")) {
            const int mx = std::atoi(s);
            n = 4;
            while (n > 1 && mx > 0 && kXYSizes[n - 1] > mx) --n;
        }
        std::vector<AxisTile> t = plan_axis(dim[d], k[d], 4, kXYSizes, n);  // multiples of 4 rows: whole row rounds in inv_xy
        if (t.empty() && n < 4) t = plan_axis(dim[d], k[d], 4, kXYSizes, 4);
        return t;
    }
    // MVD_ZMAX (tuning): largest z window.  Shorter windows make tiles small enough for a batch's
    // scratch to stay in L2 between the three kernels, at the price of more halo planes.
    int nz = 6;
    if (const char* s = std::getenv("MVD_ZMAX")) {
        const int zmax = std::atoi(s);
        while (nz > 1 && zmax > 0 && kZSizes[nz - 1] > zmax) --nz;
    }
    return plan_axis(dim[2], k[2], 1, kZSizes, nz);
}

std::unique_ptr<ConvPlan> make_plan(const int dim[3], const int k[3]) {
    auto plan = std::make_unique<ConvPlan>();
    std::vector<AxisTile> ax[3];
    for (int d = 0; d < 3; ++d) {
        plan->dim[d] = dim[d];
        plan->k[d] = k[d];
        if (dim[d] <= 0 || k[d] <= 0) throw std::invalid_argument("make_plan: non-positive size");
        ax[d] = plan_axes(dim, k, d);
        if (ax[d].empty())
            throw std::invalid_argument("make_plan: kernel size " + std::to_string(k[d]) + " in dimension " + std::to_string(d) +
                                        " exceeds the largest supported FFT window");
    }
    for (const auto& ty : ax[1])
        for (const auto& tx : ax[0])
            if (!xy_pair_supported(tx.b, ty.b))
                throw std::invalid_argument("make_plan: kernel needs 256-point FFT windows in x and y at once (" + std::to_string(k[0]) + " x " +
                                            std::to_string(k[1]) + "); a 256 x 256 plane exceeds the FFT window one SM can hold");
    std::map<std::vector<int>, size_t> index;
    for (const auto& tz : ax[2])
        for (const auto& ty : ax[1])
            for (const auto& tx : ax[0]) {
                std::vector<int> key = {tx.b, ty.b, tz.b};
                auto it = index.find(key);
                if (it == index.end()) {
                    TileClass c;
                    c.b[0] = tx.b; c.b[1] = ty.b; c.b[2] = tz.b;
                    plan->classes.push_back(std::move(c));
                    it = index.emplace(key, plan->classes.size() - 1).first;
                }
                TileClass& c = plan->classes[it->second];
                Tile t;
                t.o[0] = tx.o; t.o[1] = ty.o; t.o[2] = tz.o;
                t.e[0] = tx.e; t.e[1] = ty.e; t.e[2] = tz.e;
                t.n[0] = tx.n; t.n[1] = ty.n; t.n[2] = tz.n;
                c.tiles.push_back(t);
                c.max_nz = std::max(c.max_nz, tz.n);
            }
    for (auto& c : plan->classes) {
        MVD_CUDA(cudaMalloc(&c.d_tiles, c.tiles.size() * sizeof(Tile)));
        MVD_CUDA(cudaMemcpy(c.d_tiles, c.tiles.data(), c.tiles.size() * sizeof(Tile), cudaMemcpyHostToDevice));
        plan->max_tile_elems = std::max(plan->max_tile_elems, tile_scratch_elems(c.b[0], c.b[1], c.b[2]));
    }
    return plan;
}

KernelSpectra::~KernelSpectra() {
    for (auto* p : per_class)
        if (p) cudaFree(p);
    for (auto* p : per_class_t)
        if (p) cudaFree(p);
}

// ---- device ----------------------------------------------------------------------------
static std::mutex g_dev_mutex;
static std::map<int, std::unique_ptr<Device>> g_devices;

Device& Device::get(int id) {
    std::lock_guard<std::mutex> lk(g_dev_mutex);
    auto it = g_devices.find(id);
    if (it != g_devices.end()) {
        MVD_CUDA(cudaSetDevice(id));
        return *it->second;
    }
    int n = 0;
    MVD_CUDA(cudaGetDeviceCount(&n));
    if (id < 0 || id >= n) throw CudaError("invalid CUDA device " + std::to_string(id));
    MVD_CUDA(cudaSetDevice(id));
    std::unique_ptr<Device> d(new Device(id));
    Device& ref = *d;
    g_devices.emplace(id, std::move(d));
    return ref;
}

Device::Device(int id) : id_(id) {
    cudaDeviceProp prop;
    MVD_CUDA(cudaGetDeviceProperties(&prop, id));
    if (prop.major != 10)  // the binary holds sm_100a code only: sm_12x would fail later with an opaque error
        throw CudaError(std::string("libmvdecon is built for sm_100a (B200); device '") + prop.name + "' is sm_" +
                        std::to_string(prop.major) + std::to_string(prop.minor));
    MVD_CUDA(configure_conv_kernels());
    MVD_CUDA(configure_z_tma_kernels());
    MVD_CUDA(configure_inv_ring_kernels());
    MVD_CUDA(configure_z_warp_kernels());
    const double two_pi = 6.283185307179586476925286766559;
    for (int l = 0; l < 11; ++l) {
        const int n = 1 << l;
        std::vector<float2> h(n);
        for (int i = 0; i < n; ++i) {
            const double a = -two_pi * (double)i / (double)n;
            h[i] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
        float2* d = nullptr;
        MVD_CUDA(cudaMalloc(&d, n * sizeof(float2)));
        MVD_CUDA(cudaMemcpy(d, h.data(), n * sizeof(float2), cudaMemcpyHostToDevice));
        tw_.t[l] = d;
    }
    // Measured on B200 (cfg3): one stream and batches of a whole tile class (4.6 GB of scratch, two launches
    // per kernel and convolution) beat 256 MB batches dealt to three streams by 11 % -- bigger grids have
    // shorter tails, and concurrent kernels only fight over L2 and shared memory.
    size_t mb = std::min<size_t>(4096, prop.totalGlobalMem / (size_t)(10u << 20));
    if (const char* s = std::getenv("MVD_SCRATCH_MB")) {
        const long v = std::atol(s);
        if (v > 0) mb = (size_t)v;
    }
    scratch_budget_ = mb << 20;
    sm_count_ = prop.multiProcessorCount;
    lanes_ = 4;  // the caller's stream + three side lanes for small batches (run_convolution)
    if (const char* s = std::getenv("MVD_LANES")) {
        const int v = std::atoi(s);
        if (v > 0) lanes_ = std::min(v, (int)Scratch::kMaxLanes);
    }
    // MVD_XY_CTAS / MVD_Z_CTAS: CTAs per SM of the persistent xy / z grids (0 = one CTA per item)
    int xy = 0, z = 0;
    if (const char* s = std::getenv("MVD_XY_CTAS")) xy = std::atoi(s);
    if (const char* s = std::getenv("MVD_Z_CTAS")) z = std::atoi(s);
    set_grid_caps(xy * prop.multiProcessorCount, z * prop.multiProcessorCount);
    // MVD_L2_PERSIST=<MB> (tuning): keep batch scratch of at most that size persisting in L2 between the
    // three kernels of a batch (access policy window on the batch's stream)
    if (const char* s = std::getenv("MVD_L2_PERSIST")) {
        const long v = std::atol(s);
        if (v > 0 && prop.persistingL2CacheMaxSize > 0) {
            l2_persist_bytes_ = std::min<size_t>((size_t)v << 20, (size_t)prop.persistingL2CacheMaxSize);
            l2_window_max_ = (size_t)prop.accessPolicyMaxWindowSize;
            MVD_CUDA(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, l2_persist_bytes_));
        }
    }
}

void Scratch::init(int n) {
    if (lanes) return;
    n = std::max(1, std::min(n, (int)kMaxLanes));
    MVD_CUDA(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
    for (int i = 0; i < n; ++i) {
        MVD_CUDA(cudaStreamCreateWithFlags(&lane[i], cudaStreamNonBlocking));
        MVD_CUDA(cudaEventCreateWithFlags(&join[i], cudaEventDisableTiming));
    }
    lanes = n;
}
void Scratch::reserve(int i, size_t n) {
    if (n <= elems[i]) return;
    if (data[i]) MVD_CUDA(cudaFree(data[i]));
    data[i] = nullptr;
    elems[i] = 0;
    MVD_CUDA(cudaMalloc(&data[i], n * sizeof(float2)));
    elems[i] = n;
}
Scratch::~Scratch() {
    for (int i = 0; i < kMaxLanes; ++i) {
        if (lane[i]) { cudaStreamSynchronize(lane[i]); cudaStreamDestroy(lane[i]); }
        if (join[i]) cudaEventDestroy(join[i]);
        if (data[i]) cudaFree(data[i]);
    }
    if (fork) cudaEventDestroy(fork);
}

// ---- spectra ---------------------------------------------------------------------------
std::unique_ptr<KernelSpectra> build_spectra(Device& dev, const ConvPlan& plan, const float* kernel_host, cudaStream_t st) {
    auto spec = std::make_unique<KernelSpectra>();
    const int* k = plan.k;
    double sum = 0.0;
    for (size_t i = 0; i < (size_t)k[0] * k[1] * k[2]; ++i) sum += kernel_host[i];
    spec->kernel_sum = sum;
    // The kernel goes to the device once (K floats); padding and rotation happen there, so the host does
    // no work proportional to the window size and nothing here waits for the caller's stream.
    const size_t kn = (size_t)k[0] * k[1] * k[2];
    size_t max_elems = 0;
    for (const auto& c : plan.classes) max_elems = std::max(max_elems, (size_t)c.b[0] * c.b[1] * c.b[2]);
    float* d_k = nullptr;
    float* d_kc = nullptr;
    Tile* d_tile = nullptr;
    // stream-ordered temporaries: cudaFree would wait for the whole device, i.e. for the caller's uploads
    struct Cleanup {
        float*& a; float*& b; Tile*& t; cudaStream_t st;
        ~Cleanup() { if (a) cudaFreeAsync(a, st); if (b) cudaFreeAsync(b, st); if (t) cudaFreeAsync(t, st); }
    } cleanup{d_k, d_kc, d_tile, st};
    MVD_CUDA(cudaMallocAsync(&d_k, kn * sizeof(float), st));
    MVD_CUDA(cudaMallocAsync(&d_kc, max_elems * sizeof(float), st));
    MVD_CUDA(cudaMallocAsync(&d_tile, plan.classes.size() * sizeof(Tile), st));
    MVD_CUDA(cudaMemcpyAsync(d_k, kernel_host, kn * sizeof(float), cudaMemcpyHostToDevice, st));
    std::vector<Tile> tiles(plan.classes.size());
    for (size_t ci = 0; ci < plan.classes.size(); ++ci)
        for (int d = 0; d < 3; ++d) { tiles[ci].o[d] = 0; tiles[ci].e[d] = 0; tiles[ci].n[d] = plan.classes[ci].b[d]; }
    MVD_CUDA(cudaMemcpyAsync(d_tile, tiles.data(), tiles.size() * sizeof(Tile), cudaMemcpyHostToDevice, st));
    for (size_t ci = 0; ci < plan.classes.size(); ++ci) {
        const auto& c = plan.classes[ci];
        const int bx = c.b[0], by = c.b[1], bz = c.b[2];
        float2* d_spec = nullptr;
        MVD_CUDA(cudaMalloc(&d_spec, tile_scratch_elems(bx, by, bz) * sizeof(float2)));
        spec->per_class.push_back(d_spec);
        // kernel centre (k/2) moved to index 0, wrapped circularly (FFTConvolution.java:506-516)
        MVD_CUDA(cudaMemsetAsync(d_kc, 0, (size_t)bx * by * bz * sizeof(float), st));
        MVD_CUDA(launch_wrap_kernel(d_k, k, d_kc, c.b, st));
        Volume v;
        v.data = d_kc;
        for (int d = 0; d < 3; ++d) { v.dim[d] = c.b[d]; v.off[d] = 0; v.gmin[d] = 0; v.gdim[d] = c.b[d]; }
        v.ext = EXT_ZERO;
        v.ext_value = 0.f;
        MVD_CUDA(launch_fwd_xy(bx, by, bz, v, d_tile + ci, 1, d_spec, dev.twiddles(), st));
        // the forward x untangle omits its 1/2 for the image AND for this kernel spectrum (factor 4),
        // the unnormalised inverse contributes bx*by*bz
        const float scale = (float)(1.0 / (4.0 * (double)bx * (double)by * (double)bz));
        MVD_CUDA(launch_z_pass(bx, by, bz, 0, d_spec, nullptr, d_tile + ci, 1, scale, dev.twiddles(), st));
        g_kernel_launches += 3;
        float2* d_spec_t = nullptr;
        // 512-point classes by default; 256-point ones only with z_warp = 2 / 4 (see z_warp_covers)
        if (z_warp_supported(bz) && z_warp_covers(bz) && plane_elems(bx, by) % 16 == 0) {
            MVD_CUDA(cudaMalloc(&d_spec_t, tile_scratch_elems(bx, by, bz) * sizeof(float2)));
            spec->per_class_t.push_back(d_spec_t);
            MVD_CUDA(launch_transpose_spectrum(d_spec, d_spec_t, (int)plane_elems(bx, by), bz, st));
            ++g_kernel_launches;
        } else {
            spec->per_class_t.push_back(nullptr);
        }
    }
    MVD_CUDA(cudaStreamSynchronize(st));  // kernel_host and `tiles` may go away on return
    return spec;
}

// ---- profiler -------------------------------------------------------------------------
void Profiler::begin(int cat, double bytes, cudaStream_t st) {
    Rec r;
    r.cat = cat;
    r.bytes = bytes;
    MVD_CUDA(cudaEventCreate(&r.a));
    MVD_CUDA(cudaEventCreate(&r.b));
    MVD_CUDA(cudaEventRecord(r.a, st));
    recs.push_back(r);
}
void Profiler::end(cudaStream_t st) { MVD_CUDA(cudaEventRecord(recs.back().b, st)); }
void Profiler::collect(double ms[3], long long launches[3], double bytes[3]) {
    for (int i = 0; i < 3; ++i) { ms[i] = 0; launches[i] = 0; bytes[i] = 0; }
    for (auto& r : recs) {
        MVD_CUDA(cudaEventSynchronize(r.b));
        float t = 0.f;
        MVD_CUDA(cudaEventElapsedTime(&t, r.a, r.b));
        ms[r.cat] += t;
        launches[r.cat] += 1;
        bytes[r.cat] += r.bytes;
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    recs.clear();
}
Profiler::~Profiler() {
    for (auto& r : recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
}

bool push_split_enabled() {
    static const int v = env_flag("MVD_PUSH_SPLIT", 0);
    return v != 0;
}

void partition_for_peers(ConvPlan& plan, const PeerTarget* peers, int n) {
    if (!push_split_enabled() && n > 0) return;  // one PUSH launch over all tiles: their order does not matter
    for (TileClass& c : plan.classes) {
        auto interior = [&](const Tile& t) {
            for (int p = 0; p < n; ++p) {
                bool hit = true;
                for (int d = 0; d < 3; ++d) hit = hit && t.e[d] < peers[p].hi[d] && t.e[d] + t.n[d] > peers[p].lo[d];
                if (hit) return false;
            }
            return true;
        };
        const auto mid = std::stable_partition(c.tiles.begin(), c.tiles.end(), interior);
        c.n_interior = n > 0 ? (int)(mid - c.tiles.begin()) : -1;
        if (c.d_tiles && !c.tiles.empty())
            MVD_CUDA(cudaMemcpy(c.d_tiles, c.tiles.data(), c.tiles.size() * sizeof(Tile), cudaMemcpyHostToDevice));
    }
}

// ---- convolution -----------------------------------------------------------------------
static cudaError_t launch_inv(Device& dev, int bx, int by, int bz, int epilogue, const float2* buf, const Tile* d_tiles, int n_tiles, int zlo,
                              int max_nz, const EpilogueArgs& ep, cudaStream_t st) {
    // inv_ring: 0 = off, 1 = every epilogue, 2 = quotient / store epilogues only; MVD_RING_ORDER: 0 contiguous runs of planes per CTA, 1 round-robin
    const int ring = g_use_inv_ring.load();
    static const int ring_order = env_flag("MVD_RING_ORDER", 2);  // 2 = round-robin + L2 prefetch of the epilogue operands
    if (ring && (ring == 1 || epilogue != EPI_UPDATE) && inv_ring_supported(bx, by)) {
        const cudaError_t e = launch_inv_ring(bz, epilogue, buf, d_tiles, n_tiles, zlo, max_nz, ep, dev.sm_count(), ring_order, dev.twiddles(), st);
        if (e != cudaErrorNotSupported) return e;
    }
    return launch_inv_xy(bx, by, bz, epilogue, buf, d_tiles, n_tiles, max_nz, ep, dev.twiddles(), st);
}

void run_convolution(Device& dev, const ConvPlan& plan, const KernelSpectra& spec, const Volume& src, int epilogue,
                     const EpilogueArgs& ep, Scratch& scratch, cudaStream_t st, Profiler* prof) {
    const double epi_bytes = epilogue == EPI_STORE ? 4.0 : (epilogue == EPI_QUOTIENT ? 8.0 : (ep.weight ? 12.0 : 8.0));
    scratch.init(dev.lanes());
    // Lane 0 is the caller's stream: batches that fill the GPU run there back to back.  Batches whose
    // grids are smaller than about two waves (the 64-wide windows at the end of an axis, bricks of a
    // multi-GPU run) are dealt to the side lanes so that they overlap each other instead of each
    // paying its own launch latency and tail.  Per-kernel event timing needs everything on one stream.
    const int lanes = prof ? 1 : scratch.lanes;
    const int small_grid = 2 * 3 * dev.sm_count();
    int used = 0, next = 1;
    bool forked = false;
    for (size_t ci = 0; ci < plan.classes.size(); ++ci) {
        const TileClass& c = plan.classes[ci];
        const int bx = c.b[0], by = c.b[1], bz = c.b[2];
        const size_t te = tile_scratch_elems(bx, by, bz);
        const double pe = (double)plane_elems(bx, by);
        const int n = (int)c.tiles.size();
        int batch = (int)std::max<size_t>(1, dev.scratch_budget_bytes() / (te * sizeof(float2)));
        batch = std::min(batch, n);
        for (int s = 0; s < n; s += batch) {
            const int cnt = std::min(batch, n - s);
            const bool small = lanes > 1 && (long long)cnt * ((bz + 1) / 2) < small_grid;
            int li = 0;
            if (small) {
                li = next;
                next = next + 1 < lanes ? next + 1 : 1;
            }
            cudaStream_t ls = li ? scratch.lane[li] : st;
            if (li && !(used >> li & 1)) {
                if (!forked) {
                    MVD_CUDA(cudaEventRecord(scratch.fork, st));
                    forked = true;
                }
                MVD_CUDA(cudaStreamWaitEvent(ls, scratch.fork, 0));
                used |= 1 << li;
            }
            scratch.reserve(li, (size_t)batch * te);
            float2* buf = scratch.data[li];
            if (dev.l2_persist_bytes() > 0 && scratch.window[li] != buf) {
                const size_t bytes = (size_t)batch * te * sizeof(float2);
                if (bytes <= dev.l2_persist_bytes() && bytes <= dev.l2_window_max()) {
                    cudaStreamAttrValue av;
                    std::memset(&av, 0, sizeof(av));
                    av.accessPolicyWindow.base_ptr = buf;
                    av.accessPolicyWindow.num_bytes = bytes;
                    av.accessPolicyWindow.hitRatio = 1.f;
                    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                    MVD_CUDA(cudaStreamSetAttribute(ls, cudaStreamAttributeAccessPolicyWindow, &av));
                }
                scratch.window[li] = buf;
            }
            double nz_sum = 0, out_vox = 0;
            if (prof)
                for (int i = s; i < s + cnt; ++i) {
                    nz_sum += c.tiles[i].n[2];
                    out_vox += (double)c.tiles[i].n[0] * c.tiles[i].n[1] * c.tiles[i].n[2];
                }
            if (prof) prof->begin(0, cnt * ((double)bx * by * bz * 4.0 + pe * bz * 8.0), ls);
            MVD_CUDA(launch_fwd_xy(bx, by, bz, src, c.d_tiles + s, cnt, buf, dev.twiddles(), ls));
            if (prof) { prof->end(ls); prof->begin(1, cnt * pe * bz * 8.0 + pe * bz * 8.0 + nz_sum * pe * 8.0, ls); }
            const int zlo = c.tiles[s].e[2] - c.tiles[s].o[2];  // the same for every tile of a plan
            bool z_done = false;
            if (g_use_z_warp.load() && z_warp_supported(bz) && z_warp_covers(bz) && spec.per_class_t[ci]) {
                const cudaError_t e = launch_z_pass_warp(bx, by, bz, buf, spec.per_class_t[ci], cnt, zlo, c.max_nz, dev.sm_count(), dev.twiddles(), ls,
                                                         g_use_z_warp.load() >= 3 ? 2 : 1);
                if (e != cudaErrorNotSupported) {
                    MVD_CUDA(e);
                    z_done = true;
                }
            }
            if (z_done) {
            } else if (g_use_z_tma.load() && z_tma_supported(bz))
                MVD_CUDA(launch_z_pass_tma(bx, by, bz, buf, spec.per_class[ci], cnt, zlo, c.max_nz, dev.sm_count(),
                                           dev.twiddles(), ls));
            else
                MVD_CUDA(launch_z_pass(bx, by, bz, 1, buf, spec.per_class[ci], c.d_tiles + s, cnt, 1.f, dev.twiddles(), ls));
            if (prof) { prof->end(ls); prof->begin(2, nz_sum * pe * 8.0 + out_vox * epi_bytes, ls); }
            // MVD_PUSH_SPLIT=1: interior tiles with the plain kernel, tiles at the brick faces with the PUSH kernel (two launches
            // per batch: what paid while the push sat inside the per-voxel epilogue).  Default: one launch of the PUSH kernel --
            // with the staged push (push_plane128) it costs interior tiles nothing, and whole batches have shorter tails.
            if (push_split_enabled() && ep.n_peers > 0 && c.n_interior >= 0) {
                // interior tiles with the plain epilogue, tiles at the brick faces with the one that also
                // stores into the neighbours' ghost layers
                const int a = std::max(0, std::min(cnt, c.n_interior - s));
                EpilogueArgs plain = ep;
                plain.peers = nullptr;
                plain.n_peers = 0;
                if (a > 0) {
                    MVD_CUDA(launch_inv(dev, bx, by, bz, epilogue, buf, c.d_tiles + s, a, zlo, c.max_nz, plain, ls));
                    ++g_kernel_launches;
                }
                if (cnt - a > 0) {
                    MVD_CUDA(launch_inv(dev, bx, by, bz, epilogue, buf + (size_t)a * te, c.d_tiles + s + a, cnt - a, zlo, c.max_nz, ep, ls));
                    ++g_kernel_launches;
                }
                g_kernel_launches += 2;
            } else {
                MVD_CUDA(launch_inv(dev, bx, by, bz, epilogue, buf, c.d_tiles + s, cnt, zlo, c.max_nz, ep, ls));
                g_kernel_launches += 3;
            }
            if (prof) prof->end(ls);
        }
    }
    if (lanes > 1)
        for (int li = 1; li < lanes; ++li)
            if (used >> li & 1) {
                MVD_CUDA(cudaEventRecord(scratch.join[li], scratch.lane[li]));
                MVD_CUDA(cudaStreamWaitEvent(st, scratch.join[li], 0));
            }
}

}  // namespace mvd
