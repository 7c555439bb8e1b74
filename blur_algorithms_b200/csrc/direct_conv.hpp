// direct_conv.hpp — launcher of the fused direct-convolution blur for narrow kernels (direct_conv.cu).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace blurb200 {

constexpr int kDirectMaxPad = 20;      // widest half width the direct kernel is instantiated for (sigma <= ~6.3)

struct DirectArgs {
    const uint8_t* in;         // uint8 [rows][cols][3]
    uint8_t* out;              // uint8 [rows][cols][3] (must not alias `in`: tiles read their neighbours' pixels)
    int rows, cols, pad;
    const float* taps;         // device, [pad + 1]: tap at offset +-k (the rotated taps of getGaussian, Source.cpp:75-102)
    const float* corr_row;     // pffft Nyquist corrections (nullptr in exact mode): corr_row[ch * corr_row_pitch + row]
    int corr_row_pitch;
    const float* corr_col;     // corr_col[ch * corr_col_pitch + column]
    int corr_col_pitch;
};

// cudaErrorNotSupported when pad is outside [1, kDirectMaxPad]
cudaError_t launch_direct_gauss(const DirectArgs& a, cudaStream_t st);

}  // namespace blurb200
