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

#ifndef BLUR_OLS2_ROW_WARPS
#define BLUR_OLS2_ROW_WARPS 4
#endif
constexpr int kOls2RowWarps = BLUR_OLS2_ROW_WARPS;          // warps (independent workers) per row-pass CTA
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

// ---- bulk epilogue --------------------------------------------------------------------------
// Pass 3 with its outputs leaving through the TMA engine: once every lane holds its tile column in
// registers the tile memory is dead, so the kept outputs are written there in natural order (line a
// as floats [0, nvalid), line b behind it) and up to four lanes each issue ONE bulk shared->global
// copy (cp.async.bulk, SASS UBLKCP) per line and destination part — 2.7 KB per copy at sigma = 50
// instead of sixty-four 4-byte stores per lane.  In a sharded plan the destination is the owning
// GPU's Q over NVLink: the exchange (the distributed flip_block, Source.cpp:540) rides out of the
// FFT kernel tile by tile, in large packets, without a send buffer and without stalling the SM on
// remote store credits.  Needs 16-byte aligned pieces: V, cols, cpart and qpitch multiples of 4.
__device__ __forceinline__ void bulk_store(const void* gdst, const void* ssrc, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

constexpr int kOls2BulkB = 1056;       // float offset of line b's outputs in the tile (tile = 2112 floats)
constexpr int kOls2BulkPieces = 4;     // parts one block's outputs may span in the bulk epilogue

template <int R2, bool SHARDED>
__device__ __forceinline__ void ols2_row_pass3_bulk(cplx* __restrict__ tile, const OlsRowIO& io, const OlsRowIO::Ctx& c, int ch,
                                                    int lane, const cplx* __restrict__ w)
{
    cplx x[kOls2R1];
    const cplx* col = tile + lane;
#pragma unroll
    for (int k = 0; k < kOls2R1; ++k) x[k] = col[(R2 + 1) * k];
    __syncwarp();                                       // every lane's column is in registers: the tile is free
    CDftV<kOls2R1>::bwd_twc(x, w);
    if (io.corr) {
        const float* cr = io.corr + static_cast<size_t>(ch) * io.corr_pitch + c.ra;
        const float sg = ((c.s + lane) & 1) ? -1.f : 1.f;
        const cplx cc = cpack(sg * cr[0], c.has_b ? sg * cr[1] : 0.f);
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) x[j] = cadd(x[j], cc);
    }
    const uint32_t keep = ols2_item_mask<R2>(lane, io.ax.halo, c.nvalid);
    float* oa = reinterpret_cast<float*>(tile) + (lane - io.ax.halo);        // output i = R2 j + lane lands at float i - halo
#pragma unroll
    for (int j = 0; j < kOls2R1; ++j)
        if (keep >> j & 1u) { oa[R2 * j] = cre(x[j]); oa[kOls2BulkB + R2 * j] = cim(x[j]); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    // lane = 2 * piece + line: piece p is the run of kept columns that falls into part part0 + p (a block spans
    // at most kOls2BulkPieces parts: the launcher checks V <= (kOls2BulkPieces - 1) * cpart)
    const int line = lane & 1, piece = lane >> 1;
    if (lane < 2 * kOls2BulkPieces && (line == 0 || c.has_b)) {
        const int x_start = c.s + io.ax.halo;           // first kept column of the block
        const int nparts = SHARDED ? io.nparts : 1;
        const int part = SHARDED ? c.part0 + piece : 0;
        int x0 = x_start, x1 = x_start + c.nvalid;
        if (SHARDED) {
            x0 = max(x0, part * io.cpart);
            x1 = min(x1, (part + 1) * io.cpart);
        } else if (piece) {
            x1 = x0;
        }
        if (x1 > x0 && part < nparts) {
            float* d;
            if (SHARDED && io.peer) {
                const int chparts = io.chparts ? io.chparts : nparts;
                d = io.qpeer[part] + (static_cast<size_t>(ch * chparts + io.rank) * io.srows + c.ra + line) * io.qpitch + (x0 - part * io.cpart);
            } else {
                d = c.qa + (static_cast<size_t>(ch) * nparts + part) * io.srows * io.qpitch + static_cast<size_t>(line) * io.qpitch +
                    (x0 - part * io.cpart);
            }
            bulk_store(d, reinterpret_cast<const float*>(tile) + line * kOls2BulkB + (x0 - x_start), static_cast<uint32_t>(x1 - x0) * 4u);
        }
    }
    bulk_commit();
}

// SHARDED: the epilogue addresses the parts / peer buffers of a multi-GPU plan (io.nparts > 1)
// BULK: epilogue through bulk shared->global copies (ols2_row_pass3_bulk)
template <int R2, bool SHARDED, bool BULK>
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
    // unit = b * npairs + pair (block-major), walked with stride nw, or unit = pair * nblocks + b (pair-major),
    // every warp a contiguous run of units (a stride would pin a warp to one block index whenever nw is a
    // multiple of nblocks, and the edge blocks with their reflect fix-ups would always hit the same warps).
    // (lo, hi) = (pair, b) or (b, pair); nlo = npairs or nblocks; both orders advance without divisions.
    const int nw = gridDim.x * kOls2RowWarps;
    const bool pm = io.pair_major != 0;
    const int nblk = io.nbsel ? io.nbsel : io.ax.nblocks;        // blocks this launch transforms (OlsRowIO::bsel)
    const int nlo = pm ? nblk : io.npairs;
    const int wid = blockIdx.x * kOls2RowWarps + warp;
    const int step = pm ? 1 : nw;
    const int dlo = step % nlo, dhi = step / nlo;
    int unit = pm ? static_cast<int>(static_cast<long long>(io.nunits) * wid / nw) : wid;
    const int uend = pm ? static_cast<int>(static_cast<long long>(io.nunits) * (wid + 1) / nw) : io.nunits;
    int lo = unit % nlo, hi = unit / nlo;
    int pair = pm ? hi : lo, b = pm ? lo : hi;
    if (unit < uend) ols2_row_stage<R2>(io, pair, io.nbsel ? io.bsel[b] : b, raw, lane);
    cp_async_commit();
    __syncthreads();                                   // the multiplier table; the only block-wide barrier
    for (; unit < uend; unit += step) {
        const OlsRowIO::Ctx ctx = io.begin(pair, io.nbsel ? io.bsel[b] : b, raw, ols2_raw_pitch<R2>());
        int nlo_ = lo + dlo, nhi_ = hi + dhi;
        if (nlo_ >= nlo) { nlo_ -= nlo; ++nhi_; }
        const int npair = pm ? nhi_ : nlo_, nb = pm ? nlo_ : nhi_;
        cp_async_wait_all();
        __syncwarp();
        if (!ctx.interior) {
            for (int e = lane; e < 2 * ols2_len<R2>(); e += 32) io.fixup(ctx, e);
            __syncwarp();
        }
#pragma unroll 1
        for (int ch = 0; ch < 3; ++ch) {
            {
                cplx x[kOls2R1];
                ols2_row_pass1_compute<R2>(ctx, ch, lane, x);
                if (BULK) {                             // the previous transform's bulk copies have read the tile
                    bulk_wait_read();
                    __syncwarp();
                }
                ols2_row_pass1_store<R2>(tile, lane, x);
            }
            __syncwarp();
            if (ch == 2) {
                // the staged bytes are consumed: fetch the next unit's underneath passes 2 and 3
                if (unit + step < uend) ols2_row_stage<R2>(io, npair, io.nbsel ? io.bsel[nb] : nb, raw, lane);
                cp_async_commit();
            }
            ols2_pass2<R2, 1>(tile, lane, w, kts);
            __syncwarp();
            if (BULK)
                ols2_row_pass3_bulk<R2, SHARDED>(tile, io, ctx, ch, lane, w);
            else
                ols2_row_pass3<R2, SHARDED>(tile, io, ctx, ch, lane, w);
            // non-bulk: no barrier, the next pass 1 of this lane writes exactly the slots it has just read
        }
        pair = npair; b = nb; lo = nlo_; hi = nhi_;
    }
    if (BULK) bulk_wait_all();                          // the last copies have left shared memory and are performed
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
            const bool ok = io.qrow_hi ? (r >= io.qrow_lo && r < io.qrow_hi) : (r >= 0 && r < io.rows_total);
            // (a zero-filling copy still needs a mapped source address: the first row Q really holds)
            cp_async16_zfill(base + 4 * (R2 + 1) * j, ok ? p + j * step : io.q + static_cast<size_t>(io.qrow_hi ? io.qrow_lo : 0) * io.qpitch, ok);
        }
    }
}

__device__ __forceinline__ void cp_async_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// ---- fused interleave ---------------------------------------------------------------------------
// b g r words (4 pixels each) -> 12 interleaved bytes
__device__ __forceinline__ void ols2_il4(uint32_t b, uint32_t g, uint32_t r, uint32_t& w0, uint32_t& w1, uint32_t& w2)
{
    w0 = __byte_perm(__byte_perm(b, g, 0x1040), r, 0x3410);     // b0 g0 r0 b1
    w1 = __byte_perm(__byte_perm(g, b, 0x2601), r, 0x3250);     // g1 r1 b2 g2
    w2 = __byte_perm(__byte_perm(r, b, 0x3072), g, 0x3710);     // r2 b3 g3 r3
}

// one full 32-byte sector per lane and instruction (STG.256): a store that covers only part of a sector
// makes L2 fetch the rest from DRAM first, which is what made earlier fusions of the interleave slower
// than the separate pass
__device__ __forceinline__ void st_global_256(void* p, const uint32_t* w)
{
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]),
                 "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
}

// Item (g, b): image rows [r0, r0 + nrows) x columns [32 g, 32 g + 32) of all three channels, U -> out.
// One thread per row: twelve 8-byte loads (L2: the bytes were written moments ago by this CTA and its two
// neighbours), byte permutes, three 32-byte stores = three full sectors per row.
__device__ __forceinline__ void ols2_interleave_item(const OlsColIO& io, int g, int r0, int nrows, int tid)
{
    const int c0 = 32 * g;
    const int noct = io.noct;
    const size_t plane = static_cast<size_t>(io.rows) * 8;
    const bool fast = c0 + 32 <= io.ncols && 4 * g + 4 <= noct && (static_cast<size_t>(io.out_cols) * 3) % 32 == 0 &&
                      reinterpret_cast<uintptr_t>(io.out) % 32 == 0;
    for (int r = r0 + tid; r < r0 + nrows; r += 128) {
        if (fast) {
            uint2 v[3][4];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch)
#pragma unroll
                for (int o = 0; o < 4; ++o)
                    v[ch][o] = __ldcg(reinterpret_cast<const uint2*>(io.u + (static_cast<size_t>(ch) * noct + 4 * g + o) * plane + static_cast<size_t>(r) * 8));
            uint32_t w[24];
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                ols2_il4(v[0][o].x, v[1][o].x, v[2][o].x, w[6 * o + 0], w[6 * o + 1], w[6 * o + 2]);
                ols2_il4(v[0][o].y, v[1][o].y, v[2][o].y, w[6 * o + 3], w[6 * o + 4], w[6 * o + 5]);
            }
            uint8_t* d = io.out + (static_cast<size_t>(r) * io.out_cols + c0) * 3;
#pragma unroll
            for (int k = 0; k < 3; ++k) st_global_256(d + 32 * k, w + 8 * k);
        } else {
            for (int o = 0; o < 4 && 4 * g + o < noct; ++o)
                for (int px = 0; px < 8; ++px) {
                    const int c = c0 + 8 * o + px;
                    if (c >= io.ncols) break;
                    for (int ch = 0; ch < 3; ++ch)
                        io.out[(static_cast<size_t>(r) * io.out_cols + c) * 3 + ch] =
                            __ldcg(io.u + (static_cast<size_t>(ch) * noct + 4 * g + o) * plane + static_cast<size_t>(r) * 8 + px);
                }
        }
    }
}

// io.nunits = ols2_col_npairs(noct, nbl) (strip, block) pairs, walked in the order Ols2ColWalk::at gives
// FUSE: CTA triplets (blockIdx.x = 3 * triplet + channel) walk runs of (group, block) items, the two octet pairs
// of a group back to back; the triplet's last arriver interleaves the item (see OlsColIO::out).  gridDim.x % 3 == 0.
template <int R2, bool FUSE>
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
    int q0, q1;
    int fch = 0;                                          // FUSE: this CTA's channel
    __shared__ int s_last;
    if (FUSE) {
        // items m = g * nbl + b, nitems = nunits / 6; pair index within the CTA's run: q = 2 * m + h
        const int trip = blockIdx.x / 3, ntrip = gridDim.x / 3;
        const long long nitems = np / 6;
        fch = blockIdx.x - 3 * trip;
        q0 = 2 * static_cast<int>(nitems * trip / ntrip);
        q1 = 2 * static_cast<int>(nitems * (trip + 1) / ntrip);
    } else {
        q0 = static_cast<int>(np * blockIdx.x / gridDim.x);
        q1 = static_cast<int>(np * (blockIdx.x + 1) / gridDim.x);
    }
    // FUSE walk: pair q -> item m = q >> 1 = g * nbl + b, octet pair 2 g + (q & 1), this CTA's channel
    auto walk = [&](int q) {
        if (!FUSE) return Ols2ColWalk::at(q, io.nbl);
        const int m = q >> 1, g = m / io.nbl, b = m - g * io.nbl;
        return Ols2ColWalk{3 * (2 * g + (q & 1)) + fch, b};
    };
    Ols2ColWalk cur = walk(q0), nxt = walk(q0 + 1);
    // prologue: both members of the first pair
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        if (q0 < q1) {
            const OlsColIO::Ctx c0 = ols2_col_ctx(io, cur.strip, cur.b, m);
            if (c0.valid) ols2_col_fetch<R2>(io, c0, smem + m * kBuf + X, X, t, pol);
        }
        cp_async_commit();
    }
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
        if (FUSE && (q & 1)) {
            // this channel of item m is complete: publish, and interleave the item if the other two channels are too
            const int m = q >> 1, g = m / io.nbl, b = m - g * io.nbl;
            __threadfence();                               // this thread's U bytes are visible device-wide ...
            __syncthreads();                               // ... for every thread of the CTA
            if (tid == 0) s_last = atomicAdd(io.cnt + m, 1) == 2;
            __syncthreads();
            if (s_last) {
                __threadfence();
                const int blk = io.b0 + b;
                const int r0 = blk * io.ax.V;
                const int nrows = io.rows - r0 < io.ax.V ? io.rows - r0 : io.ax.V;
                ols2_interleave_item(io, g, r0, nrows, tid);
                if (tid == 0) io.cnt[m] = 0;               // ready for the next launch
            }
        }
        cur = nxt;
        nxt = walk(q + 2);
    }
}

}  // namespace blurb200
