// fft_row.cu — row-pass instantiation: uint8 interleaved BGR lines -> Q (Source.cpp:520-537).
#include "fft_kernels.cuh"
#include "fft_launch.hpp"

namespace blurb200 {
cudaError_t launch_fft_row(const FftGeom& g, const RowIO& io, const float2* tw, const float* kt, int batch,
                           int grid, int threads, size_t smem, cudaStream_t stream)
{
    // per-device function attribute: set on every launch (cheap), never cached process-wide
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fft_conv_kernel<RowIO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    fft_conv_kernel<RowIO><<<grid, threads, smem, stream>>>(g, io, tw, kt, batch);
    return cudaGetLastError();
}
}  // namespace blurb200
