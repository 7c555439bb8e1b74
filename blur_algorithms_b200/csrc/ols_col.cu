// ols_col.cu — column-pass instantiations of the overlap-save kernel + the TMA descriptor of Q.
#include "ols_kernels.cuh"
#include "ols_launch.hpp"

namespace blurb200 {

namespace {
using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn get_encode()
{
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeFn>(p);
    }
    return fn;
}

// Q [3][rows_total][qpitch] floats as a 3-D tensor; box = (8 columns, kOlsBoxRows rows, 1): one
// 32-byte sector per image row, landing contiguously in shared memory.
cudaError_t encode_q(CUtensorMap* map, const float* d_q, int qpitch, int rows_total)
{
    EncodeFn enc = get_encode();
    if (!enc) return cudaErrorNotSupported;
    const cuuint64_t dims[3] = {static_cast<cuuint64_t>(qpitch), static_cast<cuuint64_t>(rows_total), 3};
    const cuuint64_t strides[2] = {static_cast<cuuint64_t>(qpitch) * 4, static_cast<cuuint64_t>(qpitch) * 4 * rows_total};
    const cuuint32_t box[3] = {8, kOlsBoxRows, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(d_q), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}
}  // namespace

template <int R0>
static cudaError_t launch_col(const OlsColIO& io, const CUtensorMap& map, const float2* twA, const float2* twB,
                              const float* kt, int num_sms, cudaStream_t stream)
{
    static bool configured = false;
    constexpr int smem = ols_smem_bytes<R0, 4>();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(ols_col_kernel<R0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    const int grid = std::min(io.nunits, num_sms * ols_ctas_per_sm<R0, 4>());
    if (grid <= 0) return cudaSuccess;
    ols_col_kernel<R0><<<grid, 4 * kOlsThreads, smem, stream>>>(io, map, twA, twB, kt);
    return cudaGetLastError();
}

cudaError_t launch_ols_col(const OlsColIO& io, const float* d_q, int qpitch, int rows_total, const float2* twA,
                           const float2* twB, const float* kt, int num_sms, cudaStream_t stream)
{
    CUtensorMap map;
    cudaError_t e = encode_q(&map, d_q, qpitch, rows_total);
    if (e != cudaSuccess) return e;
    switch (io.ax.R0) {
    case 4: return launch_col<4>(io, map, twA, twB, kt, num_sms, stream);
    case 6: return launch_col<6>(io, map, twA, twB, kt, num_sms, stream);
    case 8: return launch_col<8>(io, map, twA, twB, kt, num_sms, stream);
    case 12: return launch_col<12>(io, map, twA, twB, kt, num_sms, stream);
    case 16: return launch_col<16>(io, map, twA, twB, kt, num_sms, stream);
    case 24: return launch_col<24>(io, map, twA, twB, kt, num_sms, stream);
    case 32: return launch_col<32>(io, map, twA, twB, kt, num_sms, stream);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace blurb200
