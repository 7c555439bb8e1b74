// ols_kernels.cuh — the overlap-save line-convolution kernels (sm_100a).
//
// ols_row_kernel<R0>: uint8 interleaved image rows -> Q      (Source.cpp:459-461,520-537,540)
// ols_col_kernel<R0>: Q column strips (TMA)        -> U      (Source.cpp:546-560 + rounding :567)
//
// Row pass: one CTA = 8*R0 threads = one block FFT (two image rows x L samples) at a time; the
// interleaved bytes of the two rows are staged once per unit (cp.async) and serve all three
// channels.  Column pass: one CTA = 256 threads = four block FFTs (eight image columns x L rows)
// interleaved element-wise; every thread copies the Q rows it will transform itself (cp.async,
// one 32-byte sector per image row and quad of threads) into its own tile slots as soon as it has
// finished with them in the previous unit, so the loads of unit n+1 run underneath the four
// inner passes of unit n with no staging buffer and no extra barrier.  The grid is persistent
// (gridDim.x = #SMs x resident CTAs) and strides over the units.  Resident CTAs are
// independent, so while one waits at a barrier the others keep the FMA pipe and the
// shared-memory crossbar busy.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "ols_core.cuh"

namespace blurb200 {

constexpr int kOlsThreads = 64;           // threads per transform in flight

template <int R0, int F> constexpr int ols_smem_bytes() { return 272 * R0 * 8 * F; }
template <int R0, int F> constexpr int ols_ctas_per_sm()
{
    constexpr int by_smem = (227 * 1024) / (ols_smem_bytes<R0, F>() + 1024);
    constexpr int by_warps = 8 / F;      // 16 warps of <= 128 registers
    return by_smem < 1 ? 1 : (by_smem < by_warps ? by_smem : by_warps);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---- row pass -------------------------------------------------------------------------------
// threads per CTA: pass B / mid have 16*R0 work items -> two per thread
#ifndef BLUR_OLS_ROW_T
#define BLUR_OLS_ROW_T 0      // experiment knob: force the row-pass CTA size
#endif
template <int R0> constexpr int ols_row_threads()
{
    if (BLUR_OLS_ROW_T) return BLUR_OLS_ROW_T;
    // R0 = 12: 128 threads (pass A two items each, 4 resident CTAs, ~50 KB of L1 left for the
    // twiddle tables) measured 3.5 % faster than 96 (5 CTAs): 2.13 -> 2.06 ms at 16384^2
    if (R0 == 12) return 128;
    constexpr int t = ((8 * R0 + 31) / 32) * 32;
    return t < 64 ? 64 : t;
}
template <int R0> constexpr int ols_row_smem_bytes() { return 272 * R0 * 8 + 2 * ols_raw_pitch<R0>(); }
template <int R0> constexpr int ols_row_ctas_per_sm()
{
    constexpr int by_smem = (227 * 1024) / (ols_row_smem_bytes<R0>() + 1024);
#ifdef BLUR_OLS_ROW_MINB
    constexpr int by_regs = BLUR_OLS_ROW_MINB;                // experiment knob
#else
    constexpr int by_regs = 512 / ols_row_threads<R0>();      // <= 128 registers per thread
#endif
    constexpr int c = by_smem < by_regs ? by_smem : by_regs;
    return c < 1 ? 1 : c;
}

__device__ __forceinline__ void cp_async16(void* dst, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Stage the interleaved bytes of a unit's two lines (see OlsRowIO) with 16-byte async copies.
// Vectors that would touch bytes outside [in, in_end) are copied byte by byte instead.
template <int R0, int NT>
__device__ __forceinline__ void ols_row_stage(const OlsRowIO& io, int unit, uint8_t* raw, int tid)
{
    const int pair = unit % io.npairs, b = unit / io.npairs;
    const int s = b * io.ax.V - io.ax.halo;
    const int w0 = s < 0 ? 0 : s, w1 = s + io.ax.L < io.cols ? s + io.ax.L : io.cols;
    const uint8_t* lo = io.in;
    const uint8_t* hi = io.in + static_cast<size_t>(io.rows) * io.cols * 3;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int row = 2 * pair + k;
        if (row >= io.rows) break;
        const OlsRowIO::Stage st = io.stage_of(row, s, w0, w1);
        uint8_t* dst = raw + k * ols_raw_pitch<R0>() + st.dst;
        for (int v = tid; v < st.nvec; v += NT) {
            const uint8_t* p = st.src + 16 * v;
            if (p >= lo && p + 16 <= hi) {
                cp_async16(dst + 16 * v, p);
            } else {
                for (int e = 0; e < 16; ++e)
                    if (p + e >= lo && p + e < hi) dst[16 * v + e] = p[e];
            }
        }
    }
}

template <int R0>
__global__ void __launch_bounds__(ols_row_threads<R0>(), ols_row_ctas_per_sm<R0>())
ols_row_kernel(const __grid_constant__ OlsRowIO io, const float2* __restrict__ twA_, const float2* __restrict__ twB_,
               const float* __restrict__ kt)
{
    extern __shared__ __align__(128) float2 smem_f2[];
    cplx* smem = reinterpret_cast<cplx*>(smem_f2);
    const cplx* twA = reinterpret_cast<const cplx*>(twA_);
    const cplx* twB = reinterpret_cast<const cplx*>(twB_);
    constexpr int NT = ols_row_threads<R0>();
    constexpr int NB = 16 * R0;
    uint8_t* raw = reinterpret_cast<uint8_t*>(smem + 272 * R0);
    const int tid = threadIdx.x;
    int unit = blockIdx.x;
    if (unit < io.nunits) ols_row_stage<R0, NT>(io, unit, raw, tid);
    cp_async_commit();
    for (; unit < io.nunits; unit += gridDim.x) {
        const OlsRowIO::Ctx ctx = io.begin(unit, raw, ols_raw_pitch<R0>());
        cp_async_wait_all();
        __syncthreads();
        if (!ctx.interior) {
            for (int e = tid; e < 2 * 256 * R0; e += NT) io.fixup(ctx, e);
            __syncthreads();
        }
#pragma unroll 1
        for (int ch = 0; ch < 3; ++ch) {
#pragma unroll 1
            for (int t = tid; t < 256; t += NT) ols_pass_a_fwd<R0, 1>(smem, ch, t, twA, io, ctx);
            __syncthreads();
            if (ch == 2) {
                // every thread is done with the staged bytes: fetch the next unit's underneath
                // the remaining passes of this channel
                const int next = unit + gridDim.x;
                if (next < io.nunits) ols_row_stage<R0, NT>(io, next, raw, tid);
                cp_async_commit();
            }
            // Pass B, mid and pass B' work on one k0 group (16 items) per half-warp, and item
            // `it` is the same thread in all three (NT is a multiple of 32): the data never
            // leaves the warp, so warp-level barriers are enough between them.
#pragma unroll 1
            for (int it = tid; it < NB; it += NT) ols_pass_b<false, 1>(smem, it, twB);
            __syncwarp();
#pragma unroll 1
            for (int it = tid; it < NB; it += NT) ols_mid<R0, 1>(smem, it, kt);
            __syncwarp();
#pragma unroll 1
            for (int it = tid; it < NB; it += NT) ols_pass_b<true, 1>(smem, it, twB);
            cplx w[R0];
            if (tid < 256) ols_tw_load<R0>(w, twA, tid);
            __syncthreads();
#pragma unroll 1
            for (int t = tid; t < 256; t += NT) ols_pass_a_inv<R0, 1>(smem, ch, t, t + NT, twA, w, io, ctx);
            // no barrier: the next pass A (forward) item t is this same thread on the same slots
        }
    }
}

// ---- column pass ----------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16_zfill(void* dst, const void* src, bool valid)
{
    // src-size 0 -> nothing is read and the 16 destination bytes are zero-filled; .cg = L2 only,
    // so the streamed Q data does not evict the twiddle tables from L1
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(valid ? 16 : 0) : "memory");
}

// Lanes with even X fetch, for themselves and their odd neighbour, the R0 tile rows item t owns
// for `ctx`: 16 bytes = column pairs (X, X+1) of one Q row, straight into the two adjacent slots.
template <int R0>
__device__ __forceinline__ void ols_col_fetch(const OlsColIO& io, const OlsColIO::Ctx& ctx, cplx* sx, int X, int t)
{
    if (X & 1) return;
    cplx* base = sx + 4 * ols_pad(t);
    const int r0 = ctx.s + t;
    const float* p = ctx.qa + static_cast<ptrdiff_t>(r0) * io.qpitch + 2 * X;
    const size_t step = static_cast<size_t>(256) * io.qpitch;
    if (ctx.interior) {
#pragma unroll
        for (int j = 0; j < R0; ++j) cp_async16(base + 4 * 272 * j, p + j * step);
    } else {
#pragma unroll
        for (int j = 0; j < R0; ++j) {
            const int r = r0 + 256 * j;
            const bool ok = r >= 0 && r < io.rows_total;
            cp_async16_zfill(base + 4 * 272 * j, ok ? p + j * step : io.q, ok);
        }
    }
}

template <int R0>
__global__ void __launch_bounds__(4 * kOlsThreads, ols_ctas_per_sm<R0, 4>())
ols_col_kernel(const __grid_constant__ OlsColIO io, const float2* __restrict__ twA_, const float2* __restrict__ twB_,
               const float* __restrict__ kt)
{
    extern __shared__ __align__(128) float2 smem_f2[];
    cplx* smem = reinterpret_cast<cplx*>(smem_f2);
    const cplx* twA = reinterpret_cast<const cplx*>(twA_);
    const cplx* twB = reinterpret_cast<const cplx*>(twB_);
    constexpr int F = 4, NT = F * kOlsThreads;
    constexpr int L = 256 * R0;
    const int tid = threadIdx.x;
    const int X = tid & (F - 1), sub = tid >> 2;          // work item = (transform X, index)
    cplx* sx = smem + X;
    constexpr int NB = 16 * R0;
    int unit = blockIdx.x;
    if (unit < io.nunits) {
        const OlsColIO::Ctx c0 = io.begin(unit);
#pragma unroll 1
        for (int t = sub; t < 256; t += kOlsThreads) ols_col_fetch<R0>(io, c0, sx, X, t);
    }
    cp_async_commit();
    for (; unit < io.nunits; unit += gridDim.x) {
        const OlsColIO::Ctx ctx = io.begin(unit);
        const int next = unit + gridDim.x;
        cp_async_wait_all();
        __syncwarp();                                      // rows fetched by this thread or its lane neighbour have landed
        if (!ctx.interior) {
            __syncthreads();
            for (int e = tid; e < F * L; e += NT) ols_col_patch(smem, e, ctx.s, io.ax);
            __syncthreads();
        }
#pragma unroll 1
        for (int t = sub; t < 256; t += kOlsThreads) ols_pass_a_fwd<R0, F>(sx, X, t, twA, io, ctx);
        __syncthreads();
#pragma unroll 1
        for (int it = sub; it < NB; it += kOlsThreads) ols_pass_b<false, F>(sx, it, twB);
        __syncthreads();
#pragma unroll 1
        for (int it = sub; it < NB; it += kOlsThreads) ols_mid<R0, F>(sx, it, kt);
        __syncthreads();
#pragma unroll 1
        for (int it = sub; it < NB; it += kOlsThreads) ols_pass_b<true, F>(sx, it, twB);
        cplx w[R0];
        ols_tw_load<R0>(w, twA, sub);
        __syncthreads();
        if (next < io.nunits) {
            const OlsColIO::Ctx cn = io.begin(next);
#pragma unroll 1
            for (int t = sub; t < 256; t += kOlsThreads) {
                ols_pass_a_inv<R0, F>(sx, X, t, t + kOlsThreads, twA, w, io, ctx);
                __syncwarp();                              // the odd neighbour lane has read the slots its even lane refills
                ols_col_fetch<R0>(io, cn, sx, X, t);       // same slots, same lane pair: no block barrier needed
            }
        } else {
#pragma unroll 1
            for (int t = sub; t < 256; t += kOlsThreads) ols_pass_a_inv<R0, F>(sx, X, t, t + kOlsThreads, twA, w, io, ctx);
        }
        cp_async_commit();
    }
}

}  // namespace blurb200
