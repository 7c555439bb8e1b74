// ols_col.cu — column-pass instantiations of the overlap-save kernel.
#include "ols_kernels.cuh"
#include "ols_launch.hpp"

namespace blurb200 {

template <int R0>
static cudaError_t launch_col(const OlsColIO& io, const float2* twA, const float2* twB, const float* kt, int num_sms,
                              cudaStream_t stream)
{
    constexpr int smem = ols_smem_bytes<R0, 4>();
    {   // per-device function attribute: set on every launch (cheap), never cached process-wide
        cudaError_t e = cudaFuncSetAttribute(ols_col_kernel<R0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
    }
    const int grid = std::min(io.nunits, num_sms * ols_ctas_per_sm<R0, 4>());
    if (grid <= 0) return cudaSuccess;
    ols_col_kernel<R0><<<grid, 4 * kOlsThreads, smem, stream>>>(io, twA, twB, kt);
    return cudaGetLastError();
}

cudaError_t launch_ols_col(const OlsColIO& io_in, const float* d_q, int qpitch, int rows_total, const float2* twA,
                           const float2* twB, const float* kt, int num_sms, cudaStream_t stream)
{
    OlsColIO io = io_in;
    io.q = d_q; io.qpitch = qpitch; io.rows_total = rows_total;
    switch (io.ax.R0) {
    case 4: return launch_col<4>(io, twA, twB, kt, num_sms, stream);
    case 6: return launch_col<6>(io, twA, twB, kt, num_sms, stream);
    case 8: return launch_col<8>(io, twA, twB, kt, num_sms, stream);
    case 12: return launch_col<12>(io, twA, twB, kt, num_sms, stream);
    case 16: return launch_col<16>(io, twA, twB, kt, num_sms, stream);
    case 24: return launch_col<24>(io, twA, twB, kt, num_sms, stream);
    case 32: return launch_col<32>(io, twA, twB, kt, num_sms, stream);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace blurb200
