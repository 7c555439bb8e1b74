// ols_core.cuh — overlap-save FFT convolution of image lines: per-thread building blocks.
//
// What the reference does per line (Source.cpp:520-537 / :546-560): reflect-101 pad, zero fill to
// a 2^a 3^b 5^c length N, real FFT, multiply by the kernel spectrum, inverse FFT, crop.  That is a
// LINEAR convolution of the reflect-extended line with the `ksize` Gaussian taps; the FFT length
// only has to be long enough to hold it.  Here the same convolution is evaluated block-wise
// (overlap-save): a block FFT of length L = 256*R0 consumes L consecutive samples of the
// extended line and yields V = L - 2*halo exact outputs (halo >= pad).  The result equals the
// reference's up to float rounding, but
//   * L is a small compile-time constant (R0 x 16 x 16): every shared-memory offset, twiddle
//     index and loop bound is an immediate, nothing is planned at run time;
//   * a block is 8*L bytes of shared memory (24 KB at L = 3072) instead of 8.25*N (142 KB at
//     N = 17280), so 8 independent CTAs stay resident per SM and hide each other's barriers and
//     global-memory latency;
//   * it works for any image size (no "valid size" restriction).
// Two real lines travel as one complex sequence (re = line a, im = line b); the kernel spectrum
// is real, so no split/merge step exists.  All butterflies use the packed-complex primitives of
// cplx2.cuh (FADD2 / FMUL2 / FFMA2).
//
// Decomposition of the block FFT, n = 256*n0 + 16*n1 + n2, k = k0 + R0*(k1 + 16*k2):
//   pass A  radix R0, stride 256   (reads the IO functor / writes it on the way back)
//   pass B  radix 16, stride 16
//   mid     radix 16 forward, x K[k]/L, radix 16 backward   (registers only)
//   pass B' , pass A' = the transposed passes with conjugated twiddles.
// Forward = decimation in frequency, so the spectrum is in digit-reversed order at `mid`; the
// multiplier table is stored in that order (the reference README's "unordered layout" idea).
//
// Shared-memory layout: F transforms (F = 1 or 4) are interleaved element-wise and logical element
// p of transform X lives at slot F*(p + (p >> 4)) + X: one element of padding after every 16.
// With p = 256*q + t (pass A), 256*k0 + 16*j + t (pass B) or 16*u + j (mid) the slot is
//   pass A : F*(272*q + t + (t >> 4)) + X
//   pass B : F*(272*k0 + 17*j + t) + X
//   mid    : F*(17*u + j) + X
// so every access of a work item is its base address plus a compile-time immediate, and a
// half-warp (16 consecutive work items, 8-byte elements) always covers 16 distinct bank pairs:
// conflict free in all passes.  Pass A item t reads and writes the same slot set {272*q + pad(t)},
// and the input block sits in the tile in natural order at the same slots, so pass A is in
// place per item and a thread may refill "its" slots for the next block without a barrier.
//
// All functions are host/device: tests/emul runs them on the CPU.
#pragma once
#include "dft_codelets2.cuh"

namespace blurb200 {

constexpr int kOlsMaxR0 = 32;

// Geometry of one overlap-save pass along lines of length `len`.
struct OlsAxis {
    int R0;        // first radix; 0 = axis not planned
    int L;         // block FFT length = 256 * R0
    int halo;      // context samples on each side of a block's valid outputs (>= pad)
    int V;         // valid outputs per block = L - 2*halo
    int nblocks;   // ceil(len / V)
    int len;       // line length (cols for the row pass, rows for the column pass)
    int pad;       // reflect-101 pad = (ksize - 1) / 2
    int R2;        // 0: three-factor kernels (ols_kernels.cuh); 32: two-factor kernels, L = 32*R2 (ols2_kernels.cuh)
};

// reflect-101 source index of extended-line position g, or -1 where the reference's tile holds
// zeros / nothing that reaches a kept output
BLUR_HD int ols_src(int g, int len, int pad)
{
    if (g < 0) return g >= -pad ? -g : -1;
    if (g >= len) return g <= len - 1 + pad ? 2 * (len - 1) - g : -1;
    return g;
}

BLUR_HD int ols_pad(int p) { return p + (p >> 4); }
// float2 slots of one transform's tile
BLUR_HD constexpr int ols_tile_slots(int R0) { return 272 * R0; }

// Item t of pass A owns tile indices 256 j + t.  Those inside [halo, halo + nvalid) are j in [jlo, jhi].
BLUR_HD void ols_item_range(int t, int halo, int nvalid, int& jlo, int& jhi)
{
    const int a = halo - t;                    // need 256 j >= a
    jlo = a <= 0 ? 0 : (a + 255) >> 8;
    const int b = halo + nvalid - 1 - t;       // need 256 j <= b
    jhi = b < 0 ? -1 : b >> 8;
}

// IO functor contract:
//   static constexpr bool kSmemInput      block already sits in shared memory in natural order
//   static constexpr int  kF              transforms interleaved per CTA
//   Ctx   begin(int unit) const
//   void load_item <R0>(const Ctx&, int X, int t, cplx (&x)[R0]) const    x[j] = tile_X[256 j + t]
//   void store_item<R0>(const Ctx&, int X, int t, const cplx (&x)[R0]) const
// The valid outputs of a block are tile indices [halo, halo + nvalid); item t owns j in
// [jlo, jhi] of them (ols_item_range), so the per-element test is a compare of two registers
// with an immediate.
// In all pass functions `sx` = shared-memory base of the CTA's tile + X.

// ---- pass A, forward: item (X, t), t in [0, 256)
template <int R0, int F, class IO>
BLUR_HD void ols_pass_a_fwd(cplx* sx, int X, int t, const cplx* __restrict__ twA, const IO& io,
                            const typename IO::Ctx& ctx)
{
    cplx x[R0];
    cplx* base = sx + F * ols_pad(t);
    if (IO::kSmemInput) {
#pragma unroll
        for (int j = 0; j < R0; ++j) x[j] = base[F * 272 * j];
    } else {
        io.template load_item<R0>(ctx, X, t, x);
    }
    CDft<R0, false>::run(x);
    *(base) = x[0];
#pragma unroll
    for (int q = 1; q < R0; ++q) base[F * 272 * q] = cmul(x[q], twA[(q - 1) * 256 + t]);
}

// ---- pass A, backward.  The conjugate twiddles are the first thing an item needs, and the table
// (22 KB at R0 = 12) does not stay in the few KB of L1 left beside the tiles, so each item
// consumes twiddles fetched one item ahead: `w` holds those of item t on entry and those of item
// t_next on exit (t_next >= 256: none).  The first item's are loaded before the barrier.
template <int R0>
BLUR_HD void ols_tw_load(cplx (&w)[R0], const cplx* __restrict__ twA, int t)
{
#pragma unroll
    for (int q = 1; q < R0; ++q) w[q] = twA[(q - 1) * 256 + t];
}

template <int R0, int F, class IO>
BLUR_HD void ols_pass_a_inv(const cplx* __restrict__ sx, int X, int t, int t_next, const cplx* __restrict__ twA,
                            cplx (&w)[R0], const IO& io, const typename IO::Ctx& ctx)
{
    cplx x[R0];
    const cplx* base = sx + F * ols_pad(t);
    x[0] = base[0];
#pragma unroll
    for (int q = 1; q < R0; ++q) x[q] = cmulc(base[F * 272 * q], w[q]);
    if (t_next < 256) ols_tw_load<R0>(w, twA, t_next);
    CDft<R0, true>::run(x);
    io.template store_item<R0>(ctx, X, t, x);
}

// ---- pass B: item (X, it), it in [0, 16*R0): k0 = it >> 4, t = it & 15
template <bool INV, int F>
BLUR_HD void ols_pass_b(cplx* __restrict__ sx, int it, const cplx* __restrict__ twB)
{
    const int t = it & 15;
    cplx* blk = sx + F * (272 * (it >> 4) + t);
    cplx x[16];
    if (!INV) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = blk[F * 17 * j];
        CDft<16, false>::run(x);
        *(blk) = x[0];
#pragma unroll
        for (int q = 1; q < 16; ++q) blk[F * 17 * q] = cmul(x[q], twB[(q - 1) * 16 + t]);
    } else {
        x[0] = blk[0];
#pragma unroll
        for (int q = 1; q < 16; ++q) x[q] = cmulc(blk[F * 17 * q], twB[(q - 1) * 16 + t]);
        CDft<16, true>::run(x);
#pragma unroll
        for (int j = 0; j < 16; ++j) *(blk + F * 17 * j) = x[j];
    }
}

// ---- mid: item (X, u), u in [0, 16*R0).  kt[q * (16*R0) + u] = Re K[k0 + R0*(k1 + 16 q)] / L
// with u = 16*k0 + k1.
template <int R0, int F>
BLUR_HD void ols_mid(cplx* __restrict__ sx, int u, const float* __restrict__ kt)
{
    cplx* grp = sx + F * 17 * u;
    cplx x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = grp[F * j];
    CDft<16, false>::run(x);
#pragma unroll
    for (int q = 0; q < 16; ++q) x[q] = cscale(x[q], kt[q * (16 * R0) + u]);
    CDft<16, true>::run(x);
#pragma unroll
    for (int j = 0; j < 16; ++j) *(grp + F * j) = x[j];
}

// ---- row pass IO: uint8 interleaved BGR lines -> Q -------------------------------------------
//
// HBM layouts of the overlap-save path (reference: pffft_, Source.cpp:429-570):
//   in   uint8 [rows][cols][3]
//   Q    float [3][nparts][srows][qpitch]     row-pass result, planar row-major = the reference's
//        `resf` (Source.cpp:536).  Row-pass warps write 128 contiguous bytes per store.  The
//        reference's flip_block (:540) does not exist: a column-pass CTA pulls an {8 columns x L
//        rows} strip, one 32-byte sector per image row, by asynchronous copies straight into the
//        slots its FFT passes work on (four column-pair sequences interleaved element-wise).
//        `part` = destination GPU of the column (1 on a single GPU), srows = rows per stripe,
//        qpitch = columns per part rounded up to 8.
//   U    uint8 [3][noct][rows][8]             column-pass result: per channel and column octet,
//        8 bytes per image row (rounded with +0.5f truncation, Utils.hpp:204-206)
//   out  uint8 [rows][cols][3]                interleave_oct_kernel: U -> out (replaces
//        flip_block :562 + interleave_BGR :567)
//
// unit = block * npairs + pair ; lines (2*pair, 2*pair + 1) of the stripe, all three channels.
//
// The interleaved bytes of the unit's two lines (the block's window of L pixels, 3 bytes each)
// are staged ONCE in shared memory by 128-bit asynchronous copies (cp.async; deinterleave_BGR of
// Utils.hpp:159-184 fused into the load) and then read three times, once per channel, by pass A.
// The copies of the next unit are issued as soon as the last channel has consumed the bytes and
// complete underneath that channel's remaining four passes.
//
// Staging geometry of one line: source bytes [w0*3, w1*3) of the line, w0 = max(s, 0),
// w1 = min(s + L, cols), copied in 16-byte vectors starting at the 16-byte boundary at or below
// the first byte.  The staging row (pitch 3L + 16) is addressed in TILE coordinates: byte
// (tile index i, channel c) sits at b0 + 3 i + c, with b0 in [0, 16) chosen so that source and
// destination of the copy are congruent mod 16.  For a block that overhangs the line the
// reflect-101 pixels and the zeros of the reference's tile (Source.cpp:525-529) are then written
// into the staging row by ols_row_fixup, so pass A reads every block the same way.
template <int R0> constexpr int ols_raw_pitch() { return 768 * R0 + 16; }

struct OlsRowIO {
    static constexpr bool kSmemInput = false;
    static constexpr int kF = 1;
    const uint8_t* in;     // stripe base (local row 0)
    float* q;              // Q base of this stripe
    int rows;              // valid rows in this stripe
    int cols;
    int npairs;            // ceil(rows / 2)
    int srows;             // rows per stripe (Q capacity)
    int qpitch;            // floats per Q row (columns per part, multiple of 8)
    int cpart;             // image columns per part (multiple of 8 unless nparts == 1)
    int nparts;
    int nunits;            // nblocks * npairs
    OlsAxis ax;            // len = cols
    // Fused exchange (nparts > 1, peer != 0): the epilogue stores each output straight into the
    // Q buffer of the GPU that owns its column, qpeer[part], over NVLink peer mappings, as block
    // `rank` of that buffer ([3][nparts][srows][qpitch], block index = source rank): the
    // distributed flip_block (Source.cpp:540) costs no pass over HBM and no collective call.
    // Staged exchange: every rank's part is cut into column sub-ranges ("stages"), so nparts = ranks x stages
    // virtual parts of cpart columns each, qpeer[v] = base of virtual part v's buffer, and a channel plane of
    // such a buffer holds `chparts` (= ranks) source blocks: address = qpeer[v] + ((ch * chparts + rank) * srows
    // + row) * qpitch + column within the part.  chparts = 0 means nparts (no stages).
    static constexpr int kMaxPeers = 8;
    static constexpr int kMaxParts = 64;
    float* qpeer[kMaxParts];
    int rank, peer, chparts;
    // Block selection (two-factor kernel): when nbsel > 0 the launch transforms only the overlap-save blocks
    // bsel[0 .. nbsel) of every line (a stage of the staged exchange); nunits = nbsel * npairs.
    static constexpr int kMaxSel = 64;
    int nbsel;
    unsigned char bsel[kMaxSel];
    // Order in which the persistent two-factor row grid walks its (pair, block) units: 0 = block-major
    // (unit = b * npairs + pair), 1 = pair-major (unit = pair * nblocks + b: the blocks of one row pair —
    // i.e. all destination parts of a sharded plan — are produced together, so the peer traffic of the
    // fused exchange is spread evenly over all links for the whole kernel).
    int pair_major;
    // pffft Nyquist quirk (Source.cpp:414-427) on top of the exact blocks: the reference's result is
    // exact + c * (-1)^x * A with A = alternating sum of the reflect-padded line;
    // corr[ch * corr_pitch + line] = c * A, added in the epilogue.  nullptr = exact mode.
    const float* corr;
    int corr_pitch;

    struct Stage {
        const uint8_t* src;  // 16-byte aligned global address of the first vector
        int nvec;            // 16-byte vectors to copy
        int dst;             // byte offset of the first vector in the staging row (multiple of 16)
        int b0;              // byte of (tile index 0, channel 0) in the staging row
    };
    struct Ctx {
        uint8_t* raw;        // staging area: two rows of ols_raw_pitch<R0>() bytes
        float* qa;           // Q float of (ch 0, part 0, row a, col 0)
        int ia, ib;          // byte (tile index i, channel c) of line a / b is raw[ia + 3 i + c] / raw[ib + 3 i + c]
        int ra;              // first line of the pair
        int s;               // extended-line position of tile index 0
        int w0, w1;          // staged source pixels [w0, w1)
        int nvalid;          // outputs this block keeps: tile indices [halo, halo + nvalid)
        int part0, b1;       // nparts > 1: part of the block's first output column, first column of part0 + 1
        bool has_b, interior;
    };
    BLUR_HD Stage stage_of(int row, int s, int w0, int w1) const
    {
        const uint8_t* g = in + (static_cast<size_t>(row) * cols + w0) * 3;
        const int off = static_cast<int>(reinterpret_cast<uintptr_t>(g) & 15);
        Stage st;
        st.src = g - off;
        st.nvec = (off + (w1 - w0) * 3 + 15) >> 4;
        const int lead = 3 * (w0 - s);                 // bytes of tile indices before the first staged pixel
        st.b0 = (off - lead) & 15;
        st.dst = st.b0 + lead - off;                   // multiple of 16 by construction
        return st;
    }
    BLUR_HD Ctx begin(int unit, uint8_t* raw, int raw_pitch) const { return begin(unit % npairs, unit / npairs, raw, raw_pitch); }
    BLUR_HD Ctx begin(int pair, int b, uint8_t* raw, int raw_pitch) const
    {
        Ctx c;
        c.raw = raw;
        c.ra = 2 * pair;
        c.has_b = c.ra + 1 < rows;
        c.qa = q + static_cast<size_t>(c.ra) * qpitch;
        c.s = b * ax.V - ax.halo;
        c.w0 = c.s < 0 ? 0 : c.s;
        c.w1 = c.s + ax.L < cols ? c.s + ax.L : cols;
        c.nvalid = cols - b * ax.V < ax.V ? cols - b * ax.V : ax.V;
        c.interior = c.s >= 0 && c.s + ax.L <= cols;
        c.part0 = nparts > 1 ? (b * ax.V) / cpart : 0;
        c.b1 = (c.part0 + 1) * cpart;
        c.ia = stage_of(c.ra, c.s, c.w0, c.w1).b0;
        c.ib = c.has_b ? raw_pitch + stage_of(c.ra + 1, c.s, c.w0, c.w1).b0 : c.ia;
        return c;
    }
    // Entry e of the fixup of an edge block: tile index i = e % L of line k = e / L.  Indices whose
    // extended-line position is outside [0, cols) get the reflect-101 pixel (inside the staged
    // window) or zeros.
    BLUR_HD void fixup(const Ctx& c, int e) const
    {
        const int k = e >= ax.L ? 1 : 0, i = e - k * ax.L;
        const int g = c.s + i;
        if ((g >= 0 && g < cols) || (k && !c.has_b)) return;
        uint8_t* row = c.raw + (k ? c.ib : c.ia);
        const int src = ols_src(g, cols, ax.pad);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) row[3 * i + ch] = src < 0 ? 0 : row[3 * (src - c.s) + ch];
    }
    // uint8 -> float without the conversion pipe: 0x4B000000 | b is the float 2^23 + b
    BLUR_HD static float u8f(uint8_t b)
    {
#ifdef __CUDA_ARCH__
        return __uint_as_float(0x4B000000u | b) - 8388608.f;
#else
        return static_cast<float>(b);
#endif
    }
    // X = channel
    template <int R0>
    BLUR_HD void load_item(const Ctx& c, int X, int t, cplx (&x)[R0]) const
    {
        const uint8_t* pa = c.raw + c.ia + 3 * t + X;
        const int db = c.ib - c.ia;
        if (c.has_b) {
#ifdef __CUDA_ARCH__
            // 0x4B000000 | b is the float 2^23 + b: bias both halves, remove it with ONE packed subtract
            const cplx bias = cpack(8388608.f, 8388608.f);
#pragma unroll
            for (int j = 0; j < R0; ++j)
                x[j] = csub(cpack(__uint_as_float(0x4B000000u | pa[768 * j]), __uint_as_float(0x4B000000u | pa[db + 768 * j])), bias);
#else
#pragma unroll
            for (int j = 0; j < R0; ++j) x[j] = cpack(u8f(pa[768 * j]), u8f(pa[db + 768 * j]));
#endif
        } else {
#pragma unroll
            for (int j = 0; j < R0; ++j) x[j] = cpack(u8f(pa[768 * j]), 0.f);
        }
    }
    template <int R0>
    BLUR_HD void store_item(const Ctx& c, int X, int t, const cplx (&xin)[R0]) const
    {
        cplx x[R0];
#pragma unroll
        for (int j = 0; j < R0; ++j) x[j] = xin[j];
        if (corr) {
            const float* cr = corr + static_cast<size_t>(X) * corr_pitch + c.ra;
            const float sg = ((c.s + t) & 1) ? -1.f : 1.f;       // column x0 + 256 j has the parity of x0
            const cplx cc = cpack(sg * cr[0], c.has_b ? sg * cr[1] : 0.f);
#pragma unroll
            for (int j = 0; j < R0; ++j) x[j] = cadd(x[j], cc);
        }
        int jlo, jhi;
        ols_item_range(t, ax.halo, c.nvalid, jlo, jhi);
        const int x0 = c.s + t;
        float* qc = c.qa + static_cast<size_t>(X) * nparts * srows * qpitch;
        if (nparts == 1) {
            float* d = qc + x0;
            // halo <= 256 and a full block: elements j = 1 .. R0-2 are valid for every item, only the
            // first and the last one depend on t
            const bool mid_all = ax.halo <= 256 && c.nvalid == ax.V;
            if (mid_all && c.has_b) {
                if (jlo == 0) { d[0] = cre(x[0]); d[qpitch] = cim(x[0]); }
#pragma unroll
                for (int j = 1; j < R0 - 1; ++j) { d[256 * j] = cre(x[j]); d[qpitch + 256 * j] = cim(x[j]); }
                if (jhi == R0 - 1) { d[256 * (R0 - 1)] = cre(x[R0 - 1]); d[qpitch + 256 * (R0 - 1)] = cim(x[R0 - 1]); }
            } else {
#pragma unroll
                for (int j = 0; j < R0; ++j) {
                    const bool va = j >= jlo && j <= jhi, vb = va && c.has_b;
                    if (va) d[256 * j] = cre(x[j]);
                    if (vb) d[qpitch + 256 * j] = cim(x[j]);
                }
            }
        } else if (ax.V <= 2 * cpart) {
            // The block's outputs start in part0 and span at most V <= 2*cpart columns, so an item's
            // elements fall in at most three parts: three base pointers per item, then every
            // element is a select between them plus an immediate offset (no division, no 64-bit
            // address arithmetic per element).
            const size_t rowoff = peer ? (static_cast<size_t>(X * nparts + rank) * srows + c.ra) * qpitch : 0;
            float* pp[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int part = c.part0 + k < nparts ? c.part0 + k : nparts - 1;
                float* b = peer ? qpeer[part] + rowoff : qc + static_cast<size_t>(part) * srows * qpitch;
                pp[k] = b + (static_cast<ptrdiff_t>(x0) - static_cast<ptrdiff_t>(part) * cpart);   // element x0 + 256 j lives at pp[k] + 256 j
            }
            const int e1 = c.b1 - x0, e2 = e1 + cpart;         // 256 j >= e1: part0 + 1 ; >= e2: part0 + 2
#pragma unroll
            for (int j = 0; j < R0; ++j) {
                const bool va = j >= jlo && j <= jhi, vb = va && c.has_b;
                float* d = 256 * j >= e2 ? pp[2] : (256 * j >= e1 ? pp[1] : pp[0]);
                if (va) d[256 * j] = cre(x[j]);
                if (vb) d[qpitch + 256 * j] = cim(x[j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < R0; ++j) {
                const bool va = j >= jlo && j <= jhi, vb = va && c.has_b;
                const int xx = va ? x0 + 256 * j : 0;
                const int part = xx / cpart;
                float* d = peer ? qpeer[part] + (static_cast<size_t>(X * nparts + rank) * srows + c.ra) * qpitch + (xx - part * cpart)
                                : qc + static_cast<size_t>(part) * srows * qpitch + (xx - part * cpart);
                if (va) d[0] = cre(x[j]);
                if (vb) d[qpitch] = cim(x[j]);
            }
        }
    }
};

// ---- column pass IO: Q -> U ------------------------------------------------------------------
// unit = (block * 3 + ch) * noct + oct.  One CTA transforms the four column pairs of a column
// octet at once (kF = 4).  Work item (X, t) owns tile rows 256 j + t of column pair X: it copies
// them from Q (8 bytes each, cp.async in the kernel, plain loads in the emulator; rows outside Q
// arrive as zeros), transforms them in pass A, and after the backward pass A refills the same
// slots with the next unit's rows.  Rows outside the image are patched with their reflect-101
// source (or zero) by ols_col_patch.
struct OlsColIO {
    static constexpr bool kSmemInput = true;
    static constexpr int kF = 4;
    const float* q;        // Q as received: [3][rows_total][qpitch]
    int qpitch, rows_total;
    uint8_t* u;            // U base [3][noct][rows][8]
    int rows;              // image rows = line length
    int noct;              // column octets handled here
    int nunits;
    OlsAxis ax;            // len = rows
    int b0;                // first block of this launch (the streamed host path launches block by block)
    int nbl;               // blocks in this launch (two-factor kernel: nunits = nbl * 3 * ceil(noct / 2) strip-block pairs)
    const float* corr;     // pffft Nyquist quirk, see OlsRowIO::corr: corr[ch * corr_pitch + column]
    int corr_pitch;
    // Fused interleave (two-factor kernel, ols2_col_kernel<R2, true>): CTA triplets take the three channels of the
    // same (32-column group, block) items; the CTA that finishes an item last (cnt[item] counts arrivals) reads the
    // three channels' U bytes back from L2 and writes the interleaved BGR rows of `out` (pitch out_cols pixels,
    // ncols valid columns): the separate interleave pass and its U read from DRAM disappear.
    uint8_t* out;
    int out_cols, ncols;
    int* cnt;              // one zero-initialised counter per item; the last arriver resets it
    // Row-stripe ("halo") plans of the two-factor kernel: Q and U hold only a band of image rows.  q / u are then
    // VIRTUAL bases (the address image row 0 would have), the channel planes are qplane floats / urows rows apart,
    // and only rows [qrow_lo, qrow_hi) exist in Q (outside: the reflect-101 patch or zeros, as at the image edge).
    // All four default (0) to the whole-image layout: qplane = rows_total * qpitch, urows = rows, [0, rows_total).
    long long qplane;
    int urows, qrow_lo, qrow_hi;

    struct Ctx {
        uint8_t* ua;       // U byte of (ch, oct, row 0, 0)
        const float* qa;   // Q float of (ch, row 0, column 8*oct)
        int s;             // row of tile index 0 (may be negative)
        int nvalid;        // outputs this block keeps: tile indices [halo, halo + nvalid)
        int ch, oct;
        bool interior, valid;
    };
    BLUR_HD Ctx begin(int unit) const
    {
        Ctx c;
        c.oct = unit % noct;
        const int rest = unit / noct;
        c.ch = rest % 3;
        const int b = b0 + rest / 3;
        c.valid = true;
        c.ua = u + (static_cast<size_t>(c.ch) * noct + c.oct) * rows * 8;
        c.qa = q + static_cast<size_t>(c.ch) * rows_total * qpitch + c.oct * 8;
        c.s = b * ax.V - ax.halo;
        c.nvalid = rows - b * ax.V < ax.V ? rows - b * ax.V : ax.V;
        c.interior = c.s >= 0 && c.s + ax.L <= rows;
        return c;
    }
    // source of tile row i of column pair X, or nullptr if the row is outside Q
    BLUR_HD const float* src(const Ctx& c, int X, int i) const
    {
        const int r = c.s + i;
        return (r >= 0 && r < rows_total) ? c.qa + static_cast<size_t>(r) * qpitch + 2 * X : nullptr;
    }
    // (uint8_t)(v + 0.5f): truncation, no clamp (Utils.hpp:189,204-206); negatives (undefined
    // in the reference) saturate to 0.  Returns the two bytes of a complex element packed.
    BLUR_HD static uint32_t to_u8x2(cplx v)
    {
#ifdef __CUDA_ARCH__
        const cplx t = cadd(v, cpack(0.5f, 0.5f));
        return __byte_perm(__float2uint_rz(cre(t)), __float2uint_rz(cim(t)), 0x0040);   // cvt.rzi.u32 saturates < 0 to 0
#else
        const float a = cre(v) + 0.5f, b = cim(v) + 0.5f;
        const uint32_t ua = a > 0.f ? static_cast<uint32_t>(static_cast<int>(a)) & 255u : 0u;
        const uint32_t ub = b > 0.f ? static_cast<uint32_t>(static_cast<int>(b)) & 255u : 0u;
        return ua | (ub << 8);
#endif
    }
    BLUR_HD static uint32_t to_u8(float v) { return to_u8x2(cpack(v, 0.f)) & 255u; }
    template <int R0>
    BLUR_HD void load_item(const Ctx&, int, int, cplx (&)[R0]) const {}
    template <int R0>
    BLUR_HD void store_item(const Ctx& c, int X, int t, const cplx (&xin)[R0]) const
    {
        cplx x[R0];
#pragma unroll
        for (int j = 0; j < R0; ++j) x[j] = xin[j];
        if (corr) {
            const float* cr = corr + static_cast<size_t>(c.ch) * corr_pitch + c.oct * 8 + 2 * X;
            const float sg = ((c.s + t) & 1) ? -1.f : 1.f;       // row s + t + 256 j has the parity of s + t
            const cplx cc = cpack(sg * cr[0], sg * cr[1]);
#pragma unroll
            for (int j = 0; j < R0; ++j) x[j] = cadd(x[j], cc);
        }
        int jlo, jhi;
        ols_item_range(t, ax.halo, c.nvalid, jlo, jhi);
        uint16_t* d = reinterpret_cast<uint16_t*>(c.ua + (static_cast<ptrdiff_t>(c.s + t) * 8 + 2 * X));
#pragma unroll
        for (int j = 0; j < R0; ++j)
            if (j >= jlo && j <= jhi) d[1024 * j] = static_cast<uint16_t>(to_u8x2(x[j]));
    }
};

// Patch entry e = 4*i + X of an edge block: rows outside the image take their reflect-101 source,
// which is inside the same tile, or zero.  `s` = tile base (all four transforms).
BLUR_HD void ols_col_patch(cplx* __restrict__ s, int e, int s0, const OlsAxis& ax)
{
    const int i = e >> 2, X = e & 3;
    const int g = s0 + i;
    if (g >= 0 && g < ax.len) return;
    const int src = ols_src(g, ax.len, ax.pad);
    s[4 * ols_pad(i) + X] = src < 0 ? czero() : s[4 * ols_pad(src - s0) + X];
}

// (strip, block) pairs of the two-factor column pass, see Ols2ColWalk (ols2_core.cuh)
BLUR_HD int ols2_col_npairs(int noct, int nbl) { return 2 * nbl * 3 * ((((noct + 1) / 2) + 1) / 2); }

}  // namespace blurb200
