// Streaming per-voxel kernels around the convolution engine.
#pragma once
#include <cuda_runtime.h>

namespace mvd {

static constexpr int kMaxViews = 64;
struct ViewPtrs {
    const float* img[kMaxViews];
    int n;
};

cudaError_t launch_fill(float* p, size_t n, float v, cudaStream_t st);
// FirstIteration.java:103-135: per voxel mean of the views with img > 0; accumulates the sum of
// those means (double) and the number of voxels with at least one view.
cudaError_t launch_first_iteration_sums(const ViewPtrs& views, size_t n, double* d_sum, unsigned long long* d_count, cudaStream_t st);
// MVDeconvolution.java:180-187: psi = 0 where no view has data. psi is a ghosted array.
cudaError_t launch_mask_never_updated(const ViewPtrs& views, const int brick[3], float* psi, const int psi_dim[3], const int psi_off[3],
                                      cudaStream_t st);

// kernel zero-padded to a window of size b with its centre rotated to index 0; d_out must be zeroed
cudaError_t launch_wrap_kernel(const float* d_kernel, const int k[3], float* d_out, const int b[3], cudaStream_t st);

// Neighbour-only barrier of the fused ghost exchange: thread i stores `value` into slot their_slot[i] of neighbour i's flag
// array (peer memory) and then waits until its own slot my_slot[i] holds a value >= `value` (written by that neighbour's
// barrier kernel).  The spin is bounded (~10 s): on expiry *status is set to 1 and the kernel returns.
struct PeerFlags {
    unsigned int* theirs[26];  // neighbours' flag arrays
    int their_slot[26];        // the slot of THIS brick in neighbour i's array
    int my_slot[26];           // the slot of neighbour i in this brick's array
    int n;
};
// phase 0 = signal only, 1 = wait only, 2 = signal then wait.  Do NOT launch a waiting phase ahead of the signals it waits for
// in the same stream, nor as kernels of several ranks on ONE GPU: nothing guarantees that they run at the same time.
cudaError_t launch_peer_barrier(const PeerFlags& pf, unsigned int* my_flags, unsigned int value, unsigned int* status, int phase, cudaStream_t st);

}  // namespace mvd
