// ols2_kernels.cuh — the two-factor overlap-save kernels (sm_100a), L = 32 x R2.
//
// ols2_row_kernel<R2>: uint8 interleaved image rows -> Q      (Source.cpp:459-461,520-537,540)
// ols2_col_kernel<R2>: Q column strips              -> U      (Source.cpp:546-560 + rounding :567)
//
// Row pass: every WARP is an independent worker with a private tile and a private byte-staging
// area; it walks over (block, row pair) units, stages the interleaved bytes of the two rows once
// (cp.async) and transforms the three channels one after another.  A transform is 32 items in
// each of its three passes = exactly one warp, so nothing but __syncwarp() separates the passes
// and there is no block-wide barrier in the loop at all; the warps of an SM drift apart and
// cover each other's shared-memory and global-memory phases.
//
// Column pass: one CTA of 128 threads transforms the four column pairs of a column octet
// (F = 4 interleaved tiles) so that every Q sector (32 bytes = 8 columns of one row) is fetched
// exactly once, by 16-byte cp.async straight into the slots the owning thread transforms; the
// next unit's rows are fetched as soon as pass 3 has pulled its inputs into registers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "ols2_core.cuh"
#include "ols_kernels.cuh"

namespace blurb200 {

constexpr int kOls2RowWarps = 4;          // warps (independent workers) per row-pass CTA
#ifndef BLUR_OLS2_ROW_CTAS
#define BLUR_OLS2_ROW_CTAS 3
#endif
#ifndef BLUR_OLS2_COL_CTAS
#define BLUR_OLS2_COL_CTAS 3
#endif

template <int R2> constexpr int ols2_row_warp_bytes() { return ols2_tile_slots<R2>() * 8 + 2 * ols2_raw_pitch<R2>(); }
template <int R2> constexpr int ols2_row_smem_bytes() { return ols2_len<R2>() * 4 + kOls2RowWarps * ols2_row_warp_bytes<R2>(); }
template <int R2> constexpr int ols2_col_smem_bytes() { return ols2_len<R2>() * 4 + 4 * ols2_tile_slots<R2>() * 8; }

// thread constants: w[j] = W_L^{t j} (tw2[j * 32 + t])
template <int R2>
__device__ __forceinline__ void ols2_load_twiddles(cplx (&w)[R2], const cplx* __restrict__ tw2, int t)
{
    w[0] = cpack(1.f, 0.f);
#pragma unroll
    for (int j = 1; j < R2; ++j) w[j] = tw2[j * kOls2R1 + t];
}

// Stage the interleaved bytes of a unit's two lines with 16-byte async copies issued by ONE warp.
template <int R2>
__device__ __forceinline__ void ols2_row_stage(const OlsRowIO& io, int unit, uint8_t* raw, int lane)
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
        uint8_t* dst = raw + k * ols2_raw_pitch<R2>() + st.dst;
        for (int v = lane; v < st.nvec; v += 32) {
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

template <int R2>
__global__ void __launch_bounds__(32 * kOls2RowWarps, BLUR_OLS2_ROW_CTAS)
ols2_row_kernel(const __grid_constant__ OlsRowIO io, const float2* __restrict__ tw2_, const float* __restrict__ kt)
{
    static_assert(R2 == 32, "register-resident twiddles need one item per lane in every pass");
    extern __shared__ __align__(128) float2 smem_f2[];
    float* kts = reinterpret_cast<float*>(smem_f2);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* mine = reinterpret_cast<uint8_t*>(kts + ols2_len<R2>()) + warp * ols2_row_warp_bytes<R2>();
    cplx* tile = reinterpret_cast<cplx*>(mine);
    uint8_t* raw = mine + ols2_tile_slots<R2>() * 8;
    cplx w[R2];
    ols2_load_twiddles<R2>(w, reinterpret_cast<const cplx*>(tw2_), lane);
    for (int i = threadIdx.x; i < ols2_len<R2>(); i += 32 * kOls2RowWarps) kts[i] = kt[i];
    const int nw = gridDim.x * kOls2RowWarps;
    int unit = blockIdx.x * kOls2RowWarps + warp;
    if (unit < io.nunits) ols2_row_stage<R2>(io, unit, raw, lane);
    cp_async_commit();
    __syncthreads();                                   // the multiplier table; the only block-wide barrier
    for (; unit < io.nunits; unit += nw) {
        const OlsRowIO::Ctx ctx = io.begin(unit, raw, ols2_raw_pitch<R2>());
        cp_async_wait_all();
        __syncwarp();
        if (!ctx.interior) {
            for (int e = lane; e < 2 * ols2_len<R2>(); e += 32) io.fixup(ctx, e);
            __syncwarp();
        }
#pragma unroll 1
        for (int ch = 0; ch < 3; ++ch) {
            ols2_row_pass1<R2>(tile, ctx, ch, lane);
            __syncwarp();
            if (ch == 2) {
                // the staged bytes are consumed: fetch the next unit's underneath passes 2 and 3
                const int next = unit + nw;
                if (next < io.nunits) ols2_row_stage<R2>(io, next, raw, lane);
                cp_async_commit();
            }
            ols2_pass2<R2, 1>(tile, lane, w, kts);
            __syncwarp();
            ols2_row_pass3<R2>(tile, io, ctx, ch, lane, w);
            // no barrier: the next pass 1 of this lane writes exactly the slots it has just read
        }
    }
}

// ---- column pass ------------------------------------------------------------------------------
// Lanes with even X fetch, for themselves and their odd neighbour, the 32 tile rows item n2 owns:
// 16 bytes = column pairs (X, X+1) of one Q row, straight into the two adjacent slots.
template <int R2>
__device__ __forceinline__ void ols2_col_fetch(const OlsColIO& io, const OlsColIO::Ctx& ctx, cplx* sx, int X, int n2)
{
    if (X & 1) return;
    cplx* base = sx + 4 * n2;
    const int r0 = ctx.s + n2;
    const float* p = ctx.qa + static_cast<ptrdiff_t>(r0) * io.qpitch + 2 * X;
    const size_t step = static_cast<size_t>(R2) * io.qpitch;
    if (ctx.interior) {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) cp_async16(base + 4 * (R2 + 1) * j, p + j * step);
    } else {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) {
            const int r = r0 + R2 * j;
            const bool ok = r >= 0 && r < io.rows_total;
            cp_async16_zfill(base + 4 * (R2 + 1) * j, ok ? p + j * step : io.q, ok);
        }
    }
}

template <int R2>
__global__ void __launch_bounds__(128, BLUR_OLS2_COL_CTAS)
ols2_col_kernel(const __grid_constant__ OlsColIO io, const float2* __restrict__ tw2_, const float* __restrict__ kt)
{
    static_assert(R2 == 32, "register-resident twiddles need one item per thread in every pass");
    extern __shared__ __align__(128) float2 smem_f2[];
    float* kts = reinterpret_cast<float*>(smem_f2);
    cplx* smem = reinterpret_cast<cplx*>(kts + ols2_len<R2>());
    const int tid = threadIdx.x;
    const int X = tid & 3, t = tid >> 2;                  // work item = (transform X, index t)
    cplx* sx = smem + X;
    cplx w[R2];
    ols2_load_twiddles<R2>(w, reinterpret_cast<const cplx*>(tw2_), t);
    for (int i = tid; i < ols2_len<R2>(); i += 128) kts[i] = kt[i];
    int unit = blockIdx.x;
    if (unit < io.nunits) {
        const OlsColIO::Ctx c0 = io.begin(unit);
        if (c0.valid) ols2_col_fetch<R2>(io, c0, sx, X, t);
    }
    cp_async_commit();
    __syncthreads();
    for (; unit < io.nunits; unit += gridDim.x) {
        const OlsColIO::Ctx ctx = io.begin(unit);          // ctx.valid is uniform over the CTA
        const int next = unit + gridDim.x;
        cp_async_wait_all();
        __syncwarp();                                      // rows fetched by this thread or its lane neighbour have landed
        cplx x[kOls2R1];
        if (ctx.valid) {
            if (!ctx.interior) {
                __syncthreads();
                for (int e = tid; e < 4 * ols2_len<R2>(); e += 128) ols2_col_patch<R2>(smem, e, ctx.s, io.ax);
                __syncthreads();
            }
            ols2_col_pass1<R2>(sx, t);
            __syncthreads();
            ols2_pass2<R2, 4>(sx, t, w, kts);
            __syncthreads();
            ols2_col_pass3_load<R2>(sx, t, x);
            __syncwarp();                                  // the odd neighbour has read the slots its even lane refills
        }
        if (next < io.nunits) {
            const OlsColIO::Ctx cn = io.begin(next);
            if (cn.valid) ols2_col_fetch<R2>(io, cn, sx, X, t);
        }
        cp_async_commit();
        if (ctx.valid) ols2_col_pass3_store<R2>(io, ctx, X, t, x, w);
    }
}

}  // namespace blurb200
