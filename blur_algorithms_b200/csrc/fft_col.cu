// fft_col.cu — column-pass instantiation: Q -> uint8 planar transposed U (Source.cpp:546-567).
#include "fft_kernels.cuh"
#include "fft_launch.hpp"

namespace blurb200 {
cudaError_t launch_fft_col(const FftGeom& g, const ColIO& io, const float2* tw, const float* kt, int batch,
                           int grid, int threads, size_t smem, cudaStream_t stream)
{
    // per-device function attribute: set on every launch (cheap), never cached process-wide
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fft_conv_kernel<ColIO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    fft_conv_kernel<ColIO><<<grid, threads, smem, stream>>>(g, io, tw, kt, batch);
    return cudaGetLastError();
}
}  // namespace blurb200
