// ols_row.cu — row-pass instantiations of the overlap-save kernel.
#include <algorithm>
#include <cstdlib>
#include "ols_kernels.cuh"
#include "ols_launch.hpp"

namespace blurb200 {

template <int R0>
static cudaError_t launch_row(const OlsRowIO& io, const float2* twA, const float2* twB, const float* kt, int num_sms,
                              cudaStream_t stream)
{
    constexpr int smem = ols_row_smem_bytes<R0>();
    {   // per-device function attribute: set on every launch (cheap), never cached process-wide
        cudaError_t e = cudaFuncSetAttribute(ols_row_kernel<R0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
    }
    // experiment knob: fewer resident CTAs and a smaller shared-memory carveout leave the twiddle
    // tables room in L1 (BLUR_B200_ROW_CTAS=4 BLUR_B200_ROW_CARVEOUT=86)
    static const int ctas_env = std::getenv("BLUR_B200_ROW_CTAS") ? std::atoi(std::getenv("BLUR_B200_ROW_CTAS")) : 0;
    static bool carve_done = false;
    if (!carve_done) {
        carve_done = true;
        if (const char* e = std::getenv("BLUR_B200_ROW_CARVEOUT")) {
            cudaError_t ce = cudaFuncSetAttribute(ols_row_kernel<R0>, cudaFuncAttributePreferredSharedMemoryCarveout, std::atoi(e));
            if (ce != cudaSuccess) return ce;
        }
    }
    const int per_sm = ctas_env > 0 ? std::min(ctas_env, ols_row_ctas_per_sm<R0>()) : ols_row_ctas_per_sm<R0>();
    const int grid = std::min(io.nunits, num_sms * per_sm);
    if (grid <= 0) return cudaSuccess;
    ols_row_kernel<R0><<<grid, ols_row_threads<R0>(), smem, stream>>>(io, twA, twB, kt);
    return cudaGetLastError();
}

cudaError_t launch_ols_row(const OlsRowIO& io, const float2* twA, const float2* twB, const float* kt, int num_sms,
                           cudaStream_t stream)
{
    switch (io.ax.R0) {
    case 4: return launch_row<4>(io, twA, twB, kt, num_sms, stream);
    case 6: return launch_row<6>(io, twA, twB, kt, num_sms, stream);
    case 8: return launch_row<8>(io, twA, twB, kt, num_sms, stream);
    case 12: return launch_row<12>(io, twA, twB, kt, num_sms, stream);
    case 16: return launch_row<16>(io, twA, twB, kt, num_sms, stream);
    case 24: return launch_row<24>(io, twA, twB, kt, num_sms, stream);
    case 32: return launch_row<32>(io, twA, twB, kt, num_sms, stream);
    default: return cudaErrorInvalidValue;
    }
}

int ols_resident_ctas(int R0)
{
    switch (R0) {
    case 4: return ols_row_ctas_per_sm<4>();
    case 6: return ols_row_ctas_per_sm<6>();
    case 8: return ols_row_ctas_per_sm<8>();
    case 12: return ols_row_ctas_per_sm<12>();
    case 16: return ols_row_ctas_per_sm<16>();
    case 24: return ols_row_ctas_per_sm<24>();
    case 32: return ols_row_ctas_per_sm<32>();
    default: return 0;
    }
}

}  // namespace blurb200
