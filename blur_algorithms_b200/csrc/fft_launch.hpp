// fft_launch.hpp — host-callable launchers of the fft_conv_kernel instantiations.  Each
// instantiation lives in its own translation unit (fft_row.cu / fft_col.cu) so they compile in
// parallel.
#pragma once
#include <cuda_runtime.h>
#include "blur_io.cuh"

namespace blurb200 {
cudaError_t launch_fft_row(const FftGeom& g, const RowIO& io, const float2* tw, const float* kt, int batch,
                           int grid, int threads, size_t smem, cudaStream_t stream);
cudaError_t launch_fft_col(const FftGeom& g, const ColIO& io, const float2* tw, const float* kt, int batch,
                           int grid, int threads, size_t smem, cudaStream_t stream);
}  // namespace blurb200
