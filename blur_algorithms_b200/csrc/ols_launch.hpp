// ols_launch.hpp — host-callable launchers of the overlap-save kernels (one TU per IO type so
// the instantiations compile in parallel).
#pragma once
#include <cuda_runtime.h>
#include "ols_core.cuh"

namespace blurb200 {
// grid = min(nunits, num_sms * resident CTAs); returns the CUDA error of the launch.
cudaError_t launch_ols_row(const OlsRowIO& io, const float2* twA, const float2* twB, const float* kt, int num_sms,
                           cudaStream_t stream);
// d_q: Q as received [3][rows_total][qpitch] floats; rows_total = nranks * srows
cudaError_t launch_ols_col(const OlsColIO& io, const float* d_q, int qpitch, int rows_total, const float2* twA,
                           const float2* twB, const float* kt, int num_sms, cudaStream_t stream);
int ols_resident_ctas(int R0);

// two-factor kernels (ols2_kernels.cuh); io.ax.R2 selects the instantiation.  tw2/kt: make_ols2_tables.
cudaError_t launch_ols2_row(const OlsRowIO& io, const float2* tw2, const float* kt, int num_sms, cudaStream_t stream);
cudaError_t launch_ols2_col(const OlsColIO& io, const float* d_q, int qpitch, int rows_total, const float2* tw2,
                            const float* kt, int num_sms, cudaStream_t stream);
// units (= warps) resident per SM in the row pass, for wave-aligned chunking of the sharded plan
int ols2_row_units_per_sm(int R2);
// resident CTAs per SM the two-factor kernels are compiled for (callers scale num_sms to hold a launch to fewer)
int ols2_row_ctas_per_sm();
int ols2_col_ctas_per_sm();
}  // namespace blurb200
