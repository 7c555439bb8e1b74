// ols_kernels.cuh — the overlap-save line-convolution kernels (sm_100a).
//
// ols_row_kernel<R0>: uint8 interleaved image rows -> Q      (Source.cpp:459-461,520-537,540)
// ols_col_kernel<R0>: Q column strips (TMA)        -> U      (Source.cpp:546-560 + rounding :567)
//
// Row pass: one CTA = 64 threads = one block FFT (two image rows x L samples) at a time in 8*L
// bytes of shared memory.  Column pass: one CTA = 256 threads = four block FFTs (eight image
// columns x L rows, one TMA box row = one 32-byte sector of Q) interleaved element-wise in 32*L
// bytes.  The grid is persistent (gridDim.x = #SMs x resident CTAs) and strides over the units.  Resident CTAs are independent, so while one waits at a barrier or on global
// memory the others keep the FMA pipe and the shared-memory crossbar busy.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "ols_core.cuh"

namespace blurb200 {

constexpr int kOlsThreads = 64;           // threads per transform in flight
constexpr int kOlsBoxRows = 256;          // TMA box: 256 image rows x 8 columns (32 bytes per row)

template <int R0, int F> constexpr int ols_smem_bytes() { return 256 * R0 * 8 * F + 16; }
template <int R0, int F> constexpr int ols_ctas_per_sm()
{
    constexpr int by_smem = (227 * 1024) / (ols_smem_bytes<R0, F>() + 1024);
    constexpr int by_warps = 8 / F;      // 16 warps of <= 128 registers
    return by_smem < 1 ? 1 : (by_smem < by_warps ? by_smem : by_warps);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

template <int R0>
__global__ void __launch_bounds__(kOlsThreads, ols_ctas_per_sm<R0, 1>())
ols_row_kernel(const __grid_constant__ OlsRowIO io, const float2* __restrict__ twA, const float2* __restrict__ twB,
               const float* __restrict__ kt)
{
    extern __shared__ __align__(128) float2 smem[];
    const int tid = threadIdx.x;
    constexpr int NB = 16 * R0;
    for (int unit = blockIdx.x; unit < io.nunits; unit += gridDim.x) {
        const OlsRowIO::Ctx ctx = io.begin(unit);
#pragma unroll 1
        for (int t = tid; t < 256; t += kOlsThreads) ols_pass_a_fwd<R0, 1>(smem, smem, 0, t, twA, io, ctx);
        __syncthreads();
#pragma unroll 1
        for (int it = tid; it < NB; it += kOlsThreads) ols_pass_b<false, 1>(smem, it, twB);
        __syncthreads();
#pragma unroll 1
        for (int it = tid; it < NB; it += kOlsThreads) ols_mid<R0, 1>(smem, it, kt);
        __syncthreads();
#pragma unroll 1
        for (int it = tid; it < NB; it += kOlsThreads) ols_pass_b<true, 1>(smem, it, twB);
        __syncthreads();
#pragma unroll 1
        for (int t = tid; t < 256; t += kOlsThreads) ols_pass_a_inv<R0, 1>(smem, 0, t, twA, io, ctx);
        __syncthreads();
    }
}

template <int R0>
__global__ void __launch_bounds__(4 * kOlsThreads, ols_ctas_per_sm<R0, 4>())
ols_col_kernel(const __grid_constant__ OlsColIO io, const __grid_constant__ CUtensorMap qmap,
               const float2* __restrict__ twA, const float2* __restrict__ twB, const float* __restrict__ kt)
{
    extern __shared__ __align__(128) float2 smem[];
    constexpr int F = 4, NT = F * kOlsThreads;
    constexpr int L = 256 * R0;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + F * L);
    const int tid = threadIdx.x;
    const int X = tid & (F - 1), sub = tid >> 2;          // work item = (transform X, index)
    float2* sx = smem + X;
    constexpr int NB = 16 * R0;
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;
    for (int unit = blockIdx.x; unit < io.nunits; unit += gridDim.x) {
        const OlsColIO::Ctx ctx = io.begin(unit);
        if (tid == 0) {
            // order this CTA's earlier generic-proxy accesses to the tile before the async writes
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(bar, F * L * 8);
#pragma unroll 1
            for (int k = 0; k < L; k += kOlsBoxRows)
                tma_load_3d(smem + F * k, &qmap, bar, ctx.oct * 8, ctx.s + k, ctx.ch);
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        if (!ctx.interior) {
            for (int e = tid; e < F * L; e += NT) ols_col_patch(smem, e, ctx.s, io.ax);
            __syncthreads();
        }
#pragma unroll 1
        for (int t = sub; t < 256; t += kOlsThreads) ols_pass_a_fwd<R0, F>(sx, sx, X, t, twA, io, ctx);
        __syncthreads();
#pragma unroll 1
        for (int it = sub; it < NB; it += kOlsThreads) ols_pass_b<false, F>(sx, it, twB);
        __syncthreads();
#pragma unroll 1
        for (int it = sub; it < NB; it += kOlsThreads) ols_mid<R0, F>(sx, it, kt);
        __syncthreads();
#pragma unroll 1
        for (int it = sub; it < NB; it += kOlsThreads) ols_pass_b<true, F>(sx, it, twB);
        __syncthreads();
#pragma unroll 1
        for (int t = sub; t < 256; t += kOlsThreads) ols_pass_a_inv<R0, F>(sx, X, t, twA, io, ctx);
        __syncthreads();
    }
}

}  // namespace blurb200
