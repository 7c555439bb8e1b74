// ols2.cu — instantiations and launchers of the two-factor overlap-save kernels.
#include <algorithm>
#include <cstdlib>
#include "ols2_kernels.cuh"
#include "ols_launch.hpp"

namespace blurb200 {

namespace {
// The > 48 KB dynamic shared-memory opt-in is a per-device function attribute: set it before
// every launch (it is a cheap host-side call) instead of caching a process-wide flag.
template <class K>
cudaError_t opt_in_smem(K kernel, int bytes)
{
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
}  // namespace

cudaError_t launch_ols2_row(const OlsRowIO& io, const float2* tw2, const float* kt, int num_sms, cudaStream_t stream)
{
    if (io.ax.R2 != 32) return cudaErrorInvalidValue;
    constexpr int R2 = 32;
    constexpr int smem = ols2_row_smem_bytes<R2>();
    const bool sharded = io.nparts > 1;
    // bulk epilogue: every piece a bulk copy moves must be 16-byte aligned at both ends
    static const int bulk_env = std::getenv("BLUR_B200_ROW_BULK") ? std::atoi(std::getenv("BLUR_B200_ROW_BULK")) : -1;
    const int cpart = sharded ? io.cpart : io.qpitch;
    bool bulk = io.ax.V % 4 == 0 && io.cols % 4 == 0 && io.qpitch % 4 == 0 && cpart % 4 == 0 &&
                io.ax.V <= (kOls2BulkPieces - 1) * cpart && reinterpret_cast<uintptr_t>(io.q) % 16 == 0;
    for (int k = 0; bulk && sharded && io.peer && k < io.nparts; ++k) bulk = reinterpret_cast<uintptr_t>(io.qpeer[k]) % 16 == 0;
    if (bulk_env == 0) bulk = false;
    const void* fn = sharded ? (bulk ? (const void*)ols2_row_kernel<R2, true, true> : (const void*)ols2_row_kernel<R2, true, false>)
                             : (bulk ? (const void*)ols2_row_kernel<R2, false, true> : (const void*)ols2_row_kernel<R2, false, false>);
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    static const int ctas_env = std::getenv("BLUR_B200_ROW_CTAS") ? std::atoi(std::getenv("BLUR_B200_ROW_CTAS")) : 0;
    const int per_sm = ctas_env > 0 ? std::min(ctas_env, BLUR_OLS2_ROW_CTAS) : BLUR_OLS2_ROW_CTAS;
    const int grid = std::min((io.nunits + kOls2RowWarps - 1) / kOls2RowWarps, num_sms * per_sm);
    if (grid <= 0) return cudaSuccess;
    void* args[] = {const_cast<OlsRowIO*>(&io), &tw2, &kt};
    e = cudaLaunchKernel(fn, dim3(grid), dim3(32 * kOls2RowWarps), args, smem, stream);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

cudaError_t launch_ols2_col(const OlsColIO& io_in, const float* d_q, int qpitch, int rows_total, const float2* tw2,
                            const float* kt, int num_sms, cudaStream_t stream)
{
    if (io_in.ax.R2 != 32) return cudaErrorInvalidValue;
    constexpr int R2 = 32;
    OlsColIO io = io_in;
    io.q = d_q; io.qpitch = qpitch; io.rows_total = rows_total;
    constexpr int smem = ols2_col_smem_bytes<R2>();
    const bool fuse = io.out != nullptr && io.cnt != nullptr;
    cudaError_t e = fuse ? opt_in_smem(ols2_col_kernel<R2, true>, smem) : opt_in_smem(ols2_col_kernel<R2, false>, smem);
    if (e != cudaSuccess) return e;
    static const int ctas_env = std::getenv("BLUR_B200_COL_CTAS") ? std::atoi(std::getenv("BLUR_B200_COL_CTAS")) : 0;
    const int per_sm = ctas_env > 0 ? std::min(ctas_env, BLUR_OLS2_COL_CTAS) : BLUR_OLS2_COL_CTAS;
    if (fuse) {
        // CTA triplets (one CTA per channel) over the (32-column group, block) items
        const int ntrip = std::min(io.nunits / 6, std::max(1, num_sms * per_sm / 3));
        if (ntrip <= 0) return cudaSuccess;
        ols2_col_kernel<R2, true><<<3 * ntrip, 128, smem, stream>>>(io, tw2, kt);
        return cudaGetLastError();
    }
    const int grid = std::min(io.nunits, num_sms * per_sm);       // nunits = (strip, block) pairs, dealt out in contiguous runs
    if (grid <= 0) return cudaSuccess;
    ols2_col_kernel<R2, false><<<grid, 128, smem, stream>>>(io, tw2, kt);
    return cudaGetLastError();
}

int ols2_row_units_per_sm(int R2) { return R2 == 32 ? BLUR_OLS2_ROW_CTAS * kOls2RowWarps : 0; }
int ols2_row_ctas_per_sm() { return BLUR_OLS2_ROW_CTAS; }
int ols2_col_ctas_per_sm() { return BLUR_OLS2_COL_CTAS; }

}  // namespace blurb200
