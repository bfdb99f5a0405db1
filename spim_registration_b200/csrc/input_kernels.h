// Device-side preparation of the loop's inputs (see input_kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "volume_kernels.h"

namespace mvd {

struct TransformArgs {
    double inv[12];     // row-major 3x4: world (output voxel + offset) -> view voxel coordinates
    int offset[3];      // bounding-box minimum (ProcessForDeconvolution.java:198)
    int out_dim[3];
    int src_dim[3];
    float border[3];    // Blending: pixels skipped before the blending starts
    float blending[3];  // Blending: width of the cosine ramp
    float min_value;    // MVDeconvolution.minValue (1e-4f)
};

// one view of the virtual-weight evaluation (Blending + the inverse of the view's registration)
struct BlendView {
    double inv[12];     // row-major 3x4: world voxel -> view voxel coordinates
    int dim_minus1[3];  // view size - 1 (BlendingRealRandomAccess.dimMinus1)
    float border[3];
    float blending[3];
};
struct VolumeBox {
    int offset[3];  // bounding-box minimum in world voxels
    int dim[3];
};

struct WeightPtrs {
    float* w[kMaxViews];
    int n;
};

struct PsfArgs {
    int dim[3];        // source image (extraction) or source PSF (transformation)
    int size[3];       // extracted / transformed PSF
    double inv[12];    // transformation: row-major 3x4 inverse of the view's model
    double offset[3];  // transformation: position of output voxel 0 in transformed coordinates
};

cudaError_t upload_blend_table();
cudaError_t launch_extract_psf(const float* img, const PsfArgs& a, const double* d_loc, int n_loc, float* d_psf, cudaStream_t st);
cudaError_t launch_transform_psf(const float* d_psf, const PsfArgs& a, float* d_out, cudaStream_t st);
cudaError_t launch_transform_view(const float* src, const TransformArgs& a, float* out_img, float* out_w, cudaStream_t st);
cudaError_t launch_normalize_weights(const WeightPtrs& w, size_t n, unsigned long long* d_count, int* d_min, cudaStream_t st);
cudaError_t launch_virtual_weight_sum(const BlendView* d_views, int n_views, const VolumeBox& box, float* sum_w, unsigned long long* d_count,
                                      int* d_min, cudaStream_t st);
cudaError_t launch_virtual_weight(const BlendView& view, const VolumeBox& box, const float* sum_w, double osem, float* out_w, cudaStream_t st);
cudaError_t launch_osem(float* w, size_t n, float speedup, cudaStream_t st);

}  // namespace mvd
