// ols2_core.cuh — two-factor overlap-save block FFT (L = 32 x R2): per-thread building blocks.
//
// Same job as ols_core.cuh (block-wise linear convolution of reflect-101-extended image lines
// with the Gaussian taps, Source.cpp:520-537 / :546-560), with the block FFT cut into TWO
// register-resident factors instead of three, so a transform sweeps its shared-memory tile four
// times instead of ten and loads no twiddles at all:
//
//   n = R2*n1 + n2,  k = k1 + 32*k2        (n1, k1 < 32;  n2, k2 < R2)
//   pass 1  item n2 : A[k1]  = DFT32_{n1} x[R2 n1 + n2]                       (bytes/Q -> tile)
//   pass 2  item k1 : X[k2]  = DFT_R2_{n2}( A[k1][n2] * W_L^{n2 k1} )
//                     C[n2]  = IDFT_R2_{k2}( X[k2] * Re K[k1 + 32 k2] / L )    (tile -> tile)
//   pass 3  item n2 : y[R2 n1 + n2] = IDFT32_{k1}( C[k1][n2] * W_L^{-n2 k1} )  (tile -> Q/U)
//
// The inter-pass twiddles W_L^{jt} ride into the first butterflies of the pass-2 / pass-3
// codelets (gen_codelets3.py: a + w*b as two FMAs, a - w*b = 2a - s as one).  With R2 = 32 the
// twiddles an item needs are W_L^{t j}, j = 1..31, where t is the item index in BOTH passes —
// and an item index belongs to one thread for the whole kernel, so the 31 twiddles are thread
// constants held in registers (loaded once per kernel).  The multiplier Re K/L comes from a 4 KB
// table in shared memory.
//
// Tile layout: element (k1, n2) lives at slot F*(k1*(R2+1) + n2) + X (X = transform of F
// interleaved ones).  Pass 1/3 item n2 owns slots {j*(R2+1) + n2}, pass 2 item k1 owns
// {k1*(R2+1) + j}: every pass is in place per item, every access is base + immediate, and
// consecutive items hit consecutive (pass 1/3) or stride-(R2+1) (pass 2) slots: conflict free.
// The input block sits in the tile in natural order at slot (i / R2)*(R2+1) + i % R2.
//
// All functions are host/device: tests/emul runs them on the CPU.
#pragma once
#include "dft_codelets3.cuh"
#include "ols_core.cuh"

namespace blurb200 {

constexpr int kOls2R1 = 32;

template <int R2> BLUR_HD constexpr int ols2_pitch() { return R2 + 1; }
template <int R2> BLUR_HD constexpr int ols2_tile_slots() { return kOls2R1 * (R2 + 1); }
template <int R2> BLUR_HD constexpr int ols2_len() { return kOls2R1 * R2; }
template <int R2> BLUR_HD constexpr int ols2_raw_pitch() { return 3 * kOls2R1 * R2 + 16; }
// slot of tile index i (natural order)
template <int R2> BLUR_HD int ols2_slot(int i) { return (i / R2) * (R2 + 1) + (i % R2); }

// Item t of pass 1/3 owns tile indices R2 j + t.  Those inside [halo, halo + nvalid) are j in [jlo, jhi].
template <int R2>
BLUR_HD void ols2_item_range(int t, int halo, int nvalid, int& jlo, int& jhi)
{
    const int a = halo - t;
    jlo = a <= 0 ? 0 : (a + R2 - 1) / R2;
    const int b = halo + nvalid - 1 - t;
    jhi = b < 0 ? -1 : b / R2;
}

// the same as a bit mask: bit j set <=> output j of item t is kept
template <int R2>
BLUR_HD uint32_t ols2_item_mask(int t, int halo, int nvalid)
{
    int jlo, jhi;
    ols2_item_range<R2>(t, halo, nvalid, jlo, jhi);
    if (jhi < jlo) return 0u;
    if (jhi > 31) jhi = 31;
    return ((2u << jhi) - 1u) & ~((1u << jlo) - 1u);
}

// ---- pass 2 (both kernels): item k1 of transform X; sx = tile base + X
template <int R2, int F>
BLUR_HD void ols2_pass2(cplx* __restrict__ sx, int k1, const cplx* __restrict__ w, const float* __restrict__ kt)
{
    cplx* row = sx + F * (R2 + 1) * k1;
    cplx x[R2];
#pragma unroll
    for (int j = 0; j < R2; ++j) x[j] = row[F * j];
    CDftV<R2>::fwd_tw(x, w);
#pragma unroll
    for (int q = 0; q < R2; ++q) x[q] = cscale(x[q], kt[q * kOls2R1 + k1]);
    CDftV<R2>::bwd(x);
#pragma unroll
    for (int j = 0; j < R2; ++j) row[F * j] = x[j];
}

// ---- row pass ---------------------------------------------------------------------------------
// pass 1: item n2 reads the staged interleaved bytes of the unit's two lines (OlsRowIO::Ctx),
// channel ch, and leaves A[k1][n2] in the tile.
template <int R2>
BLUR_HD void ols2_row_pass1_compute(const OlsRowIO::Ctx& c, int ch, int n2, cplx (&x)[kOls2R1])
{
    const uint8_t* pa = c.raw + c.ia + 3 * n2 + ch;
    const int db = c.ib - c.ia;
#ifdef __CUDA_ARCH__
    // uint8 -> float without the conversion pipe and without a subtraction per element: one PRMT
    // drops the byte into bits 8..15 of 0x47000000, i.e. the float 32768 + b.  The DFT sees
    // b + 32768; every butterfly of the radix-32 codelet only adds or subtracts bias-carrying
    // values with each other before any multiplication, all of them integers below 2^21, so the
    // arithmetic is exact: the bias cancels in every difference and arrives as 32 * 32768 = 2^20
    // in output 0, where ONE subtraction per item removes it.
    if (c.has_b) {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j)
            x[j] = cpack(__uint_as_float(__byte_perm(pa[3 * R2 * j], 0x47000000u, 0x7604)),
                         __uint_as_float(__byte_perm(pa[db + 3 * R2 * j], 0x47000000u, 0x7604)));
    } else {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) x[j] = cpack(__uint_as_float(__byte_perm(pa[3 * R2 * j], 0x47000000u, 0x7604)), 0.f);
    }
    CDftV<kOls2R1>::fwd(x);
    x[0] = csub(x[0], cpack(1048576.f, c.has_b ? 1048576.f : 0.f));
#else
    if (c.has_b) {
        for (int j = 0; j < kOls2R1; ++j) x[j] = cpack(OlsRowIO::u8f(pa[3 * R2 * j]), OlsRowIO::u8f(pa[db + 3 * R2 * j]));
    } else {
        for (int j = 0; j < kOls2R1; ++j) x[j] = cpack(OlsRowIO::u8f(pa[3 * R2 * j]), 0.f);
    }
    CDftV<kOls2R1>::fwd(x);
#endif
}

template <int R2>
BLUR_HD void ols2_row_pass1_store(cplx* __restrict__ tile, int n2, const cplx (&x)[kOls2R1])
{
    cplx* col = tile + n2;
#pragma unroll
    for (int k = 0; k < kOls2R1; ++k) col[(R2 + 1) * k] = x[k];
}

template <int R2>
BLUR_HD void ols2_row_pass1(cplx* __restrict__ tile, const OlsRowIO::Ctx& c, int ch, int n2)
{
    cplx x[kOls2R1];
    ols2_row_pass1_compute<R2>(c, ch, n2, x);
    ols2_row_pass1_store<R2>(tile, n2, x);
}

// pass 3: item n2 -> Q.  Output j of the item is image column s + R2 j + n2 of lines (ra, ra+1).
// SHARDED = false: single part (nparts == 1), the only epilogue the single-GPU kernel carries.
template <int R2, bool SHARDED>
BLUR_HD void ols2_row_pass3(const cplx* __restrict__ tile, const OlsRowIO& io, const OlsRowIO::Ctx& c, int ch, int n2,
                            const cplx* __restrict__ w)
{
    cplx x[kOls2R1];
    const cplx* col = tile + n2;
#pragma unroll
    for (int k = 0; k < kOls2R1; ++k) x[k] = col[(R2 + 1) * k];
    CDftV<kOls2R1>::bwd_twc(x, w);
    if (io.corr) {
        const float* cr = io.corr + static_cast<size_t>(ch) * io.corr_pitch + c.ra;
        const float sg = ((c.s + n2) & 1) ? -1.f : 1.f;       // R2 is even: column x0 + R2 j has the parity of x0
        const cplx cc = cpack(sg * cr[0], c.has_b ? sg * cr[1] : 0.f);
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) x[j] = cadd(x[j], cc);
    }
    const uint32_t keep = ols2_item_mask<R2>(n2, io.ax.halo, c.nvalid);
    const int x0 = c.s + n2;
    const int nparts = SHARDED ? io.nparts : 1, qpitch = io.qpitch;
    const int chparts = SHARDED && io.chparts ? io.chparts : nparts;       // source blocks per channel plane of a peer buffer
    float* qc = c.qa + static_cast<size_t>(ch) * nparts * io.srows * qpitch;
    if (!SHARDED) {
        float* d = qc + x0;
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j)
            if (keep >> j & 1u) d[R2 * j] = cre(x[j]);
        if (c.has_b) {
            d += qpitch;
#pragma unroll
            for (int j = 0; j < kOls2R1; ++j)
                if (keep >> j & 1u) d[R2 * j] = cim(x[j]);
        }
    } else if (io.ax.V <= 2 * io.cpart) {
        // the block's outputs span at most three parts: three base pointers, then a select per element
        const size_t rowoff = io.peer ? (static_cast<size_t>(ch * chparts + io.rank) * io.srows + c.ra) * qpitch : 0;
        float* pp[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int part = c.part0 + k < nparts ? c.part0 + k : nparts - 1;
            float* b = io.peer ? io.qpeer[part] + rowoff : qc + static_cast<size_t>(part) * io.srows * qpitch;
            pp[k] = b + (static_cast<ptrdiff_t>(x0) - static_cast<ptrdiff_t>(part) * io.cpart);
        }
        const int e1 = c.b1 - x0, e2 = e1 + io.cpart;
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) {
            const bool va = keep >> j & 1u, vb = va && c.has_b;
            float* d = R2 * j >= e2 ? pp[2] : (R2 * j >= e1 ? pp[1] : pp[0]);
            if (va) d[R2 * j] = cre(x[j]);
            if (vb) d[qpitch + R2 * j] = cim(x[j]);
        }
    } else {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) {
            const bool va = keep >> j & 1u, vb = va && c.has_b;
            const int xx = va ? x0 + R2 * j : 0;
            const int part = xx / io.cpart;
            float* d = io.peer ? io.qpeer[part] + (static_cast<size_t>(ch * chparts + io.rank) * io.srows + c.ra) * qpitch + (xx - part * io.cpart)
                               : qc + static_cast<size_t>(part) * io.srows * qpitch + (xx - part * io.cpart);
            if (va) d[0] = cre(x[j]);
            if (vb) d[qpitch] = cim(x[j]);
        }
    }
}

// (uint8_t)(v + 0.5f) of both halves (Utils.hpp:189,204-206: float add, then truncation).  The blur of
// non-negative data with non-negative taps is >= -rounding noise, so v + 0.5f > 0 and truncation
// = floor, which one packed add in round-down mode delivers in the low mantissa byte of
// 2^23 + t: no float->int conversion instruction (those run on the quarter-rate XU pipe).
// CLAMP (pffft-quirk plans: the Nyquist correction can drive values below -0.5): negatives -> 0, the
// saturation OlsColIO::to_u8x2 documents.
template <bool CLAMP>
BLUR_HD uint32_t ols2_to_u8x2(cplx v)
{
#ifdef __CUDA_ARCH__
    cplx t = cadd(v, cpack(0.5f, 0.5f));
    if (CLAMP) t = cpack(fmaxf(cre(t), 0.f), fmaxf(cim(t), 0.f));
    const cplx m = cpack(8388608.f, 8388608.f);
    asm("add.rm.f32x2 %0, %1, %2;" : "=l"(t) : "l"(t), "l"(m));
    return __byte_perm(__float_as_uint(cre(t)), __float_as_uint(cim(t)), 0x0040);
#else
    return OlsColIO::to_u8x2(v);
#endif
}

// ---- column pass ------------------------------------------------------------------------------
// Four column pairs (one column octet) are interleaved element-wise (F = 4); sx = tile base + X.
// pass 1: item n2 transforms the Q rows it (or its lane neighbour) copied into its own slots.
template <int R2>
BLUR_HD void ols2_col_pass1(cplx* __restrict__ sx, int n2)
{
    cplx x[kOls2R1];
    cplx* col = sx + 4 * n2;
#pragma unroll
    for (int j = 0; j < kOls2R1; ++j) x[j] = col[4 * (R2 + 1) * j];
    CDftV<kOls2R1>::fwd(x);
#pragma unroll
    for (int k = 0; k < kOls2R1; ++k) col[4 * (R2 + 1) * k] = x[k];
}

// pass 3, first half: the item's tile column into registers (its slots are free afterwards)
template <int R2>
BLUR_HD void ols2_col_pass3_load(const cplx* __restrict__ sx, int n2, cplx (&x)[kOls2R1])
{
    const cplx* col = sx + 4 * n2;
#pragma unroll
    for (int k = 0; k < kOls2R1; ++k) x[k] = col[4 * (R2 + 1) * k];
}

// pass 3, second half: inverse DFT, rounding, U bytes.  Output j is image row s + R2 j + n2.
template <int R2>
BLUR_HD void ols2_col_pass3_store(const OlsColIO& io, const OlsColIO::Ctx& c, int X, int n2, cplx (&x)[kOls2R1],
                                  const cplx* __restrict__ w)
{
    CDftV<kOls2R1>::bwd_twc(x, w);
    if (io.corr) {
        const float* cr = io.corr + static_cast<size_t>(c.ch) * io.corr_pitch + c.oct * 8 + 2 * X;
        const float sg = ((c.s + n2) & 1) ? -1.f : 1.f;
        const cplx cc = cpack(sg * cr[0], sg * cr[1]);
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j) x[j] = cadd(x[j], cc);
    }
    const uint32_t keep = ols2_item_mask<R2>(n2, io.ax.halo, c.nvalid);
    uint16_t* d = reinterpret_cast<uint16_t*>(c.ua + (static_cast<ptrdiff_t>(c.s + n2) * 8 + 2 * X));
    if (io.corr) {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j)
            if (keep >> j & 1u) d[4 * R2 * j] = static_cast<uint16_t>(ols2_to_u8x2<true>(x[j]));
    } else {
#pragma unroll
        for (int j = 0; j < kOls2R1; ++j)
            if (keep >> j & 1u) d[4 * R2 * j] = static_cast<uint16_t>(ols2_to_u8x2<false>(x[j]));
    }
}

// Pair q of the column pass -> (strip, block).  A strip is (octet pair j, channel): 16 float columns =
// 64 bytes of a Q row.  DRAM hands L2 whole 128-byte lines, and at this kernel's fill rate a line lives
// in L2 for only a few tens of microseconds, so the two pairs that share a line (j = 2g, 2g+1) are walked
// back to back, then the next block of the same strips (its halo rows are still resident), then the
// channels, then the next line group.  io.nunits = 2 * nbl * 3 * ceil(ceil(noct / 2) / 2).
struct Ols2ColWalk {
    int strip, b;
    BLUR_HD static Ols2ColWalk at(int q, int nbl)
    {
        const int h = q & 1, r = q >> 1;
        const int b = r % nbl, rest = r / nbl;
        const int ch = rest % 3, g = rest / 3;
        return Ols2ColWalk{3 * (2 * g + h) + ch, b};
    }
};

// sub-unit context of (strip, b, member m): channel = strip % 3, octet = 2 * (strip / 3) + m
BLUR_HD OlsColIO::Ctx ols2_col_ctx(const OlsColIO& io, int strip, int b, int m)
{
    OlsColIO::Ctx c;
    c.ch = strip % 3;
    c.oct = 2 * (strip / 3) + m;
    c.valid = c.oct < io.noct;
    if (!c.valid) c.oct = io.noct - 1;
    const int blk = io.b0 + b;
    c.ua = io.u + (static_cast<size_t>(c.ch) * io.noct + c.oct) * (io.urows ? io.urows : io.rows) * 8;
    c.qa = io.q + static_cast<size_t>(c.ch) * (io.qplane ? static_cast<size_t>(io.qplane) : static_cast<size_t>(io.rows_total) * io.qpitch) + c.oct * 8;
    c.s = blk * io.ax.V - io.ax.halo;
    c.nvalid = io.rows - blk * io.ax.V < io.ax.V ? io.rows - blk * io.ax.V : io.ax.V;
    c.interior = c.s >= 0 && c.s + io.ax.L <= io.rows;
    return c;
}

// Patch entry e = 4*i + X of an edge block (rows outside the image take their reflect-101 source,
// which is inside the same tile, or zero).  `s` = tile base (all four transforms).
template <int R2>
BLUR_HD void ols2_col_patch(cplx* __restrict__ s, int e, int s0, const OlsAxis& ax)
{
    const int i = e >> 2, X = e & 3;
    const int g = s0 + i;
    if (g >= 0 && g < ax.len) return;
    const int src = ols_src(g, ax.len, ax.pad);
    s[4 * ols2_slot<R2>(i) + X] = src < 0 ? czero() : s[4 * ols2_slot<R2>(src - s0) + X];
}

}  // namespace blurb200
