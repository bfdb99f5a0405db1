// Host side of the tiled convolution engine: tile planner, per-device resources,
// kernel-spectrum construction and the batched launch sequence.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "conv_kernels.h"
#include "conv_types.h"

namespace mvd {

struct CudaError : std::runtime_error {
    explicit CudaError(const std::string& s) : std::runtime_error(s) {}
};
void cuda_check(cudaError_t e, const char* what);
#define MVD_CUDA(x) ::mvd::cuda_check((x), #x)

extern std::atomic<long long> g_kernel_launches;
// run-time switches between kernel generations (A/B measurements; mvdSetTuning / environment MVD_Z_TMA, MVD_INV_TMA)
extern std::atomic<int> g_use_z_tma;
extern std::atomic<int> g_use_inv_ring;
extern std::atomic<int> g_use_z_warp;

// ---- planner ---------------------------------------------------------------------------
struct AxisTile {
    int b;  // FFT length of the window
    int o;  // window origin (brick coordinates, may be negative)
    int e;  // first output voxel
    int n;  // number of output voxels
};
// Cover [0,n) with overlap-save windows for a kernel of length k whose centre sits at k/2.
// Outputs of a window start `lo` voxels after its origin, lo = roundup(k-1-k/2, align).
std::vector<AxisTile> plan_axis(int n, int k, int align, const int* sizes, int num_sizes);

struct TileClass {
    int b[3];
    std::vector<Tile> tiles;
    Tile* d_tiles = nullptr;  // device copy
    int max_nz = 0;
    // fused ghost exchange: tiles [0, n_interior) touch no neighbour's ghost layers and run the plain
    // epilogue, the rest runs the PUSH epilogue (partition_for_peers); -1 = not partitioned
    int n_interior = -1;
};

struct ConvPlan {
    int dim[3];
    int k[3];
    std::vector<TileClass> classes;
    size_t max_tile_elems = 0;  // largest per-tile scratch (complex elements)
    ~ConvPlan();
    ConvPlan() = default;
    ConvPlan(const ConvPlan&) = delete;
    ConvPlan& operator=(const ConvPlan&) = delete;
};
// Builds the plan and uploads the tile lists to the current device.
std::unique_ptr<ConvPlan> make_plan(const int dim[3], const int k[3]);
bool same_plan(const ConvPlan& p, const int dim[3], const int k[3]);
// Orders the tiles of every class interior first (stable) with respect to the neighbours' boxes and
// uploads them again; n = 0 removes the partition.  No kernel using the plan may be in flight.
void partition_for_peers(ConvPlan& plan, const PeerTarget* peers, int n);
bool push_split_enabled();  // MVD_PUSH_SPLIT=1: interior / boundary tiles in separate launches (run_convolution)

// Spectra of one kernel for every tile class of a plan (resident in HBM).
struct KernelSpectra {
    std::vector<float2*> per_class;
    std::vector<float2*> per_class_t;  // transposed copy [position][z] for the warp-local z pass (512-point classes), else nullptr
    double kernel_sum = 0.0;
    ~KernelSpectra();
    KernelSpectra() = default;
    KernelSpectra(const KernelSpectra&) = delete;
    KernelSpectra& operator=(const KernelSpectra&) = delete;
};

// ---- per-device resources --------------------------------------------------------------
class Device {
public:
    static Device& get(int id);  // throws CudaError
    int id() const { return id_; }
    const TwiddleTables& twiddles() const { return tw_; }
    size_t scratch_budget_bytes() const { return scratch_budget_; }
    int lanes() const { return lanes_; }
    int sm_count() const { return sm_count_; }
    size_t l2_persist_bytes() const { return l2_persist_bytes_; }
    size_t l2_window_max() const { return l2_window_max_; }
    std::mutex legacy_mutex;  // serialises the drop-in symbol per device

private:
    explicit Device(int id);
    int id_;
    TwiddleTables tw_{};
    size_t scratch_budget_;
    int lanes_ = 1;
    int sm_count_ = 148;
    size_t l2_persist_bytes_ = 0, l2_window_max_ = 0;
};

// Scratch for the tile batches of one convolution stream.  Lane 0 is the caller's stream; batches too
// small to fill the GPU are dealt round-robin to the side lanes 1 .. lanes-1 (each with its own
// scratch buffer), which fork from / join into the caller's stream with events.
struct Scratch {
    static constexpr int kMaxLanes = 4;
    float2* data[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
    size_t elems[kMaxLanes] = {0, 0, 0, 0};
    cudaStream_t lane[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t fork = nullptr;
    cudaEvent_t join[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
    int lanes = 0;  // 0 = not initialised
    const float2* window[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};  // buffer the lane's L2 access-policy window covers
    void init(int n);
    void reserve(int lane_index, size_t n);
    ~Scratch();
    Scratch() = default;
    Scratch(const Scratch&) = delete;
    Scratch& operator=(const Scratch&) = delete;
};

// Optional per-kernel timing with CUDA events on the launching stream (bench.py roofline).
// Categories: 0 fwd_xy, 1 z_pass, 2 inv_xy.  `bytes` are the ALGORITHMIC bytes of the launch
// (DESIGN.md "byte model"): each pass counts its input once and its output once, the z pass
// additionally the kernel spectrum once per launch, inv_xy the arrays of its fused epilogue.
struct Profiler {
    struct Rec { int cat; cudaEvent_t a, b; double bytes; };
    std::vector<Rec> recs;
    void begin(int cat, double bytes, cudaStream_t st);
    void end(cudaStream_t st);
    // synchronises the events and accumulates; clears the records
    void collect(double ms[3], long long launches[3], double bytes[3]);
    ~Profiler();
};

std::unique_ptr<KernelSpectra> build_spectra(Device& dev, const ConvPlan& plan, const float* kernel_host, cudaStream_t st);

// out-of-place tiled convolution of `src`; the epilogue decides what is stored.
void run_convolution(Device& dev, const ConvPlan& plan, const KernelSpectra& spec, const Volume& src, int epilogue,
                     const EpilogueArgs& ep, Scratch& scratch, cudaStream_t st, Profiler* prof = nullptr);

}  // namespace mvd
