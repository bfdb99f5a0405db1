// Launchers of the tiled FFT-convolution kernels (conv_kernels.cu).
#pragma once
#include "conv_types.h"

namespace mvd {

struct Twiddles;  // fft_device.cuh

struct TwiddleTables {
    const float2* t[11];
};

// true when (bx,by) / bz are sizes the kernels are instantiated for
bool xy_size_supported(int b);
bool y_size_supported(int b);
// windows with a 256-point side exist for kernels too wide for 128 (one 512-thread CTA per SM); 256 x 256 planes do not fit one SM
bool xy_pair_supported(int bx, int by);
bool z_size_supported(int b);

size_t plane_elems(int bx, int by);                 // (bx/2+1)*by complex per z-plane
size_t tile_scratch_elems(int bx, int by, int bz);  // complex elements per tile

cudaError_t launch_fwd_xy(int bx, int by, int bz, const Volume& src, const Tile* d_tiles, int n_tiles, float2* scratch,
                          const TwiddleTables& tw, cudaStream_t st);
// mode 0: forward only + scale (kernel spectrum); mode 1: forward * kspec -> inverse
cudaError_t launch_z_pass(int bx, int by, int bz, int mode, float2* scratch, const float2* kspec, const Tile* d_tiles, int n_tiles,
                          float scale, const TwiddleTables& tw, cudaStream_t st);
cudaError_t launch_inv_xy(int bx, int by, int bz, int epilogue, const float2* scratch, const Tile* d_tiles, int n_tiles, int max_nz,
                          const EpilogueArgs& ep, const TwiddleTables& tw, cudaStream_t st);
// z pass with TMA data movement (z_tma_kernels.cu): bulk tensor loads / stores of whole work items, two thread groups and
// three shared-memory slots per CTA.  mode 1 only; bz in {128, 256, 512}.
bool z_tma_supported(int bz);
cudaError_t configure_z_tma_kernels();
cudaError_t launch_z_pass_tma(int bx, int by, int bz, float2* scratch, const float2* kspec, int n_tiles, int zlo, int max_nz, int num_sms,
                              const TwiddleTables& tw, cudaStream_t st);
// z pass with warp-local lines (z_warp_kernels.cu): 512-point windows, needs the transposed kernel spectrum ([position][z])
bool z_warp_supported(int bz);
cudaError_t configure_z_warp_kernels();
cudaError_t launch_transpose_spectrum(const float2* spec, float2* spec_t, int pe, int bz, cudaStream_t st);
cudaError_t launch_z_pass_warp(int bx, int by, int bz, float2* scratch, const float2* kspec_t, int n_tiles, int zlo, int max_nz, int num_sms,
                               const TwiddleTables& tw, cudaStream_t st, int generation = 1);
// inverse xy pass for 128 x 128 windows fed by bulk copies into a shared-memory ring (inv_ring_kernels.cu)
bool inv_ring_supported(int bx, int by);
cudaError_t configure_inv_ring_kernels();
// zlo = first plane of a tile that is handed on; cudaErrorNotSupported: use launch_inv_xy
cudaError_t launch_inv_ring(int bz, int epilogue, const float2* scratch, const Tile* d_tiles, int n_tiles, int zlo, int max_nz,
                            const EpilogueArgs& ep, int num_sms, int interleave, const TwiddleTables& tw, cudaStream_t st);
void set_grid_caps(int xy_cap, int z_cap);  // persistent-grid sizes, 0 = one CTA per work item
cudaError_t configure_conv_kernels();  // opt in to > 48 KB dynamic shared memory

}  // namespace mvd
