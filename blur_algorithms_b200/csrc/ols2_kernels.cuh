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
template <int R2> constexpr int ols2_col_smem_bytes() { return ols2_len<R2>() * 4 + 2 * 4 * ols2_tile_slots<R2>() * 8; }

// thread constants: w[j] = W_L^{t j} (tw2[j * 32 + t])
template <int R2>
__device__ __forceinline__ void ols2_load_twiddles(cplx (&w)[R2], const cplx* __restrict__ tw2, int t)
{
    w[0] = cpack(1.f, 0.f);
#pragma unroll
    for (int j = 1; j < R2; ++j) w[j] = tw2[j * kOls2R1 + t];
}

// Stage the interleaved bytes of the two lines of unit (pair, b) with 16-byte async copies issued by
// ONE warp.  Vectors that would touch bytes outside [in, in_end) are copied byte by byte instead
// (only the first / last line of a stripe can need that).
template <int R2>
__device__ __forceinline__ void ols2_row_stage(const OlsRowIO& io, int pair, int b, uint8_t* raw, int lane)
{
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
        if (st.src >= lo && st.src + 16 * st.nvec <= hi) {
            for (int v = lane; v < st.nvec; v += 32) cp_async16(dst + 16 * v, st.src + 16 * v);
        } else {
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
}

// SHARDED: the epilogue addresses the parts / peer buffers of a multi-GPU plan (io.nparts > 1)
template <int R2, bool SHARDED>
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
    // unit = b * npairs + pair, walked with stride nw: (pair, b) advance without divisions
    const int nw = gridDim.x * kOls2RowWarps, npairs = io.npairs;
    const int dpair = nw % npairs, db = nw / npairs;
    int unit = blockIdx.x * kOls2RowWarps + warp;
    int pair = unit % npairs, b = unit / npairs;
    if (unit < io.nunits) ols2_row_stage<R2>(io, pair, b, raw, lane);
    cp_async_commit();
    __syncthreads();                                   // the multiplier table; the only block-wide barrier
    for (; unit < io.nunits; unit += nw) {
        const OlsRowIO::Ctx ctx = io.begin(pair, b, raw, ols2_raw_pitch<R2>());
        int npair = pair + dpair, nb = b + db;
        if (npair >= npairs) { npair -= npairs; ++nb; }
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
                if (unit + nw < io.nunits) ols2_row_stage<R2>(io, npair, nb, raw, lane);
                cp_async_commit();
            }
            ols2_pass2<R2, 1>(tile, lane, w, kts);
            __syncwarp();
            ols2_row_pass3<R2, SHARDED>(tile, io, ctx, ch, lane, w);
            // no barrier: the next pass 1 of this lane writes exactly the slots it has just read
        }
        pair = npair; b = nb;
    }
}

// ---- column pass ------------------------------------------------------------------------------
// A CTA owns a contiguous run of (strip, block) pairs, strip = (column-octet PAIR, channel), walked
// block after block down the strip.  The two octets of the pair live in two tile buffers that are
// transformed alternately: while one is in its FFT passes the other one's rows for the next block
// are in flight (a whole unit of prefetch distance), the two 32-byte halves of every 64-byte run
// of a Q row are requested a few microseconds apart by the same SM, and so are the 2*halo rows that
// vertically adjacent blocks share: each reaches DRAM once.
// cp.async flavour of the Q strip fetch (experiment knob): 0 plain .cg, 1 .cg with
// L2::128B prefetch, 2 .cg with an evict_last L2 policy
#ifndef BLUR_OLS2_COL_HINT
#define BLUR_OLS2_COL_HINT 0
#endif
__device__ __forceinline__ void cp_async16_q(void* dst, const void* src, uint64_t pol)
{
#if BLUR_OLS2_COL_HINT == 1
    asm volatile("cp.async.cg.shared.global.L2::128B [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
#elif BLUR_OLS2_COL_HINT == 2
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "l"(pol) : "memory");
#else
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
#endif
}
__device__ __forceinline__ uint64_t l2_policy_evict_last()
{
    uint64_t pol = 0;
#if BLUR_OLS2_COL_HINT == 2
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
#endif
    return pol;
}

// Lanes with even X fetch, for themselves and their odd neighbour, the 32 tile rows item n2 owns:
// 16 bytes = column pairs (X, X+1) of one Q row, straight into the two adjacent slots.
template <int R2>
__device__ __forceinline__ void ols2_col_fetch(const OlsColIO& io, const OlsColIO::Ctx& ctx, cplx* sx, int X, int n2, uint64_t pol)
{
    if (X & 1) return;
    cplx* base = sx + 4 * n2;
    const int r0 = ctx.s + n2;
    const float* p = ctx.qa + static_cast<ptrdiff_t>(r0) * io.qpitch + 2 * X;
    const size_t step = static_cast<size_t>(R2) * io.qpitch;
    if (ctx.interior) {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) cp_async16_q(base + 4 * (R2 + 1) * j, p + j * step, pol);
    } else {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) {
            const int r = r0 + R2 * j;
            const bool ok = r >= 0 && r < io.rows_total;
            cp_async16_zfill(base + 4 * (R2 + 1) * j, ok ? p + j * step : io.q, ok);
        }
    }
}

__device__ __forceinline__ void cp_async_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// io.nunits = number of (strip, block) pairs = nstrips * io.nbl, nstrips = 3 * ceil(noct / 2)
template <int R2>
__global__ void __launch_bounds__(128, BLUR_OLS2_COL_CTAS)
ols2_col_kernel(const __grid_constant__ OlsColIO io, const float2* __restrict__ tw2_, const float* __restrict__ kt)
{
    static_assert(R2 == 32, "register-resident twiddles need one item per thread in every pass");
    extern __shared__ __align__(128) float2 smem_f2[];
    float* kts = reinterpret_cast<float*>(smem_f2);
    cplx* smem = reinterpret_cast<cplx*>(kts + ols2_len<R2>());
    constexpr int kBuf = 4 * ols2_tile_slots<R2>();       // slots of one buffer (four interleaved tiles)
    const int tid = threadIdx.x;
    const int X = tid & 3, t = tid >> 2;                  // work item = (transform X, index t)
    cplx w[R2];
    ols2_load_twiddles<R2>(w, reinterpret_cast<const cplx*>(tw2_), t);
    for (int i = tid; i < ols2_len<R2>(); i += 128) kts[i] = kt[i];
    const uint64_t pol = l2_policy_evict_last();
    // this CTA's run of pairs
    const long long np = io.nunits;
    const int q0 = static_cast<int>(np * blockIdx.x / gridDim.x), q1 = static_cast<int>(np * (blockIdx.x + 1) / gridDim.x);
    Ols2ColWalk cur{q0 / io.nbl, q0 % io.nbl}, nxt = cur;
    // prologue: both members of the first pair
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        if (q0 < q1) {
            const OlsColIO::Ctx c0 = ols2_col_ctx(io, cur.strip, cur.b, m);
            if (c0.valid) ols2_col_fetch<R2>(io, c0, smem + m * kBuf + X, X, t, pol);
        }
        cp_async_commit();
    }
    nxt.next(io.nbl);
    __syncthreads();
    for (int q = q0; q < q1; ++q) {
#pragma unroll 1
        for (int m = 0; m < 2; ++m) {
            cplx* buf = smem + m * kBuf;
            cplx* sx = buf + X;
            const OlsColIO::Ctx ctx = ols2_col_ctx(io, cur.strip, cur.b, m);      // ctx.valid is uniform over the CTA
            cp_async_wait_1();                             // this buffer's rows have landed (the other buffer's may be in flight)
            __syncwarp();                                  // ... including those fetched by the lane neighbour
            cplx x[kOls2R1];
            if (ctx.valid) {
                if (!ctx.interior) {
                    __syncthreads();
                    for (int e = tid; e < 4 * ols2_len<R2>(); e += 128) ols2_col_patch<R2>(buf, e, ctx.s, io.ax);
                    __syncthreads();
                }
                ols2_col_pass1<R2>(sx, t);
                __syncthreads();
                ols2_pass2<R2, 4>(sx, t, w, kts);
                __syncthreads();
                ols2_col_pass3_load<R2>(sx, t, x);
                __syncwarp();                              // the odd neighbour has read the slots its even lane refills
            }
            if (q + 1 < q1) {
                const OlsColIO::Ctx cn = ols2_col_ctx(io, nxt.strip, nxt.b, m);
                if (cn.valid) ols2_col_fetch<R2>(io, cn, sx, X, t, pol);
            }
            cp_async_commit();
            if (ctx.valid) ols2_col_pass3_store<R2>(io, ctx, X, t, x, w);
        }
        cur = nxt;
        nxt.next(io.nbl);
    }
}

}  // namespace blurb200
