// Shared host/device descriptors of the tiled FFT-convolution engine.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mvd {

enum ExtMode : int { EXT_ZERO = 0, EXT_CONST = 1, EXT_MIRROR_SINGLE = 2, EXT_CLAMP = 3 };

// Out-of-bounds rule of the *source* of a convolution.  The reference feeds
// conv1 with Views.extendMirrorSingle(psi) (MVDeconFFTThreads.java:23,57) and
// conv2 with Views.extendValue(quotient, 1.0f) (MVDeconFFTThreads.java:42,82;
// MVDeconFFT.java:490).  Zero extension serves the drop-in block symbol and
// the kernel preparation (Views.extendZero, MVDeconFFT.java:201-227).  EXT_CLAMP repeats the last
// pixel ("extendlastpixel" of the separable-convolution hook, CUDASeparableConvolution.java:12).
struct Volume {
    const float* data;   // x fastest, index = x + dim[0]*(y + dim[1]*z), local array
    int dim[3];          // array extent (brick plus ghost layers)
    int off[3];          // array index of brick voxel 0 (= low ghost width)
    int gmin[3];         // global coordinate of brick voxel 0
    int gdim[3];         // global domain size the extension rule refers to
    int ext;             // ExtMode
    float ext_value;     // EXT_CONST fill
};

// One overlap-save tile.  All coordinates are brick coordinates (brick voxel 0 = 0).
struct Tile {
    int o[3];   // origin of the B^3 input window (may be negative / exceed the array)
    int e[3];   // origin of the effective (output) region
    int n[3];   // size of the effective region
};

// Epilogue selection for the inverse xy pass.
enum Epilogue : int {
    EPI_STORE = 0,     // out = conv
    EPI_QUOTIENT = 1,  // out = img > 0 ? img / conv : 1         (MVDeconvolution.java:473-525)
    EPI_UPDATE = 2     // psi = update(psi, conv, weight, lambda) (MVDeconvolution.java:582-703)
};

// Ghost layers of a neighbouring brick that this brick's epilogue fills directly through peer
// memory (NVLink stores, or plain stores when the neighbour lives on the same device): every
// output voxel inside [lo, hi) (brick coordinates) is also written to
// base[((z + shift[2]) * dim[1] + (y + shift[1])) * dim[0] + x + shift[0]].
struct PeerTarget {
    float* base;   // the neighbour's ghosted psi / quotient array
    int dim[3];    // its extent
    int shift[3];  // brick coordinate -> index in the neighbour's array
    int lo[3], hi[3];
};
constexpr int kMaxPeers = 26;

struct EpilogueArgs {
    float* out;            // EPI_STORE / EPI_QUOTIENT target, EPI_UPDATE: psi (read + write)
    const float* img;      // EPI_QUOTIENT: observed view
    const float* weight;   // EPI_UPDATE: per-voxel weight or nullptr
    float weight_const;    // EPI_UPDATE: used when weight == nullptr
    float add_const;       // added to the convolution result before the epilogue
    double lambda;         // Tikhonov parameter (0 = off)
    float tik_hi, tik_lo;  // 2 lambda as a float pair (make_tikhonov), filled by the host: left to the kernel, the compiler
                           // recomputed the pair per voxel (2 DADD, 3 F2F: 5 % of the update kernel's instructions)
    float min_value;       // 1e-4f in the reference (MVDeconvolution.java:49)
    double* sum_change;    // EPI_UPDATE statistics (device scalars) or nullptr
    unsigned int* max_change_bits;
    int out_dim[3];        // extent of the out array (brick plus ghost layers)
    int out_off[3];        // array index of brick voxel 0 inside out
    int aux_dim[3];        // extent of img / weight (brick only, no ghost)
    const PeerTarget* peers;  // device array of n_peers neighbours whose ghost layers are pushed, or nullptr
    int n_peers;
};

}  // namespace mvd
