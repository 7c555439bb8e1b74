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
// Shared-memory layout: logical element p lives at slot p ^ ((p >> 4) & 15).  The XOR only
// touches the low four bits, so
//   pass A : slot = 256*q + (t ^ (t>>4 & 15))            one XOR per item, immediates after that
//   pass B : slot = 256*k0 + 16*j + (t ^ j)
//   mid    : slot = 16*u + (j ^ (u & 15))
// and in every pass the 16 lanes of a half-warp hit 16 distinct 8-byte bank pairs: conflict
// free without padding.  A block that arrives by TMA in NATURAL order (column pass) is consumed
// by pass A in place: item t reads slots 256*j + t and writes 256*q + swz(t); swz(t) stays inside
// t's aligned group of 16 items, i.e. inside the same half-warp of the same round, whose loads
// all precede its stores.
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
};

// reflect-101 source index of extended-line position g, or -1 where the reference's tile holds
// zeros / nothing that reaches a kept output
BLUR_HD int ols_src(int g, int len, int pad)
{
    if (g < 0) return g >= -pad ? -g : -1;
    if (g >= len) return g <= len - 1 + pad ? 2 * (len - 1) - g : -1;
    return g;
}

BLUR_HD int ols_swz16(int t) { return t ^ ((t >> 4) & 15); }

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
//   Ctx   begin(int unit) const
//   void load_item <R0>(const Ctx&, int t, float2 (&x)[R0]) const    x[j] = tile[256 j + t]
//   void store_item<R0>(const Ctx&, int t, const float2 (&x)[R0]) const
// The valid outputs of a block are tile indices [halo, halo + nvalid); item t owns j in
// [jlo, jhi] of them (ols_item_range), so the per-element test is a compare of two registers
// with an immediate.

// ---- pass A, forward: item t in [0, 256)
template <int R0, class IO>
BLUR_HD void ols_pass_a_fwd(float2* __restrict__ s, int t, const float2* __restrict__ twA, const IO& io,
                            const typename IO::Ctx& ctx)
{
    float2 x[R0];
    if (IO::kSmemInput) {
#pragma unroll
        for (int j = 0; j < R0; ++j) x[j] = s[256 * j + t];
    } else {
        io.template load_item<R0>(ctx, t, x);
    }
    CDft<R0, false>::run(x);
    const int ta = ols_swz16(t);
    s[ta] = x[0];
#pragma unroll
    for (int q = 1; q < R0; ++q) s[256 * q + ta] = cmul(x[q], twA[(q - 1) * 256 + t]);
}

// ---- pass A, backward: item t in [0, 256)
template <int R0, class IO>
BLUR_HD void ols_pass_a_inv(const float2* __restrict__ s, int t, const float2* __restrict__ twA, const IO& io,
                            const typename IO::Ctx& ctx)
{
    float2 x[R0];
    const int ta = ols_swz16(t);
    x[0] = s[ta];
#pragma unroll
    for (int q = 1; q < R0; ++q) x[q] = cmulc(s[256 * q + ta], twA[(q - 1) * 256 + t]);
    CDft<R0, true>::run(x);
    io.template store_item<R0>(ctx, t, x);
}

// ---- pass B: item it in [0, 16*R0): k0 = it >> 4, t = it & 15
template <bool INV>
BLUR_HD void ols_pass_b(float2* __restrict__ s, int it, const float2* __restrict__ twB)
{
    const int t = it & 15;
    float2* blk = s + ((it >> 4) << 8);
    float2 x[16];
    if (!INV) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = blk[16 * j + (t ^ j)];
        CDft<16, false>::run(x);
        blk[t] = x[0];
#pragma unroll
        for (int q = 1; q < 16; ++q) blk[16 * q + (t ^ q)] = cmul(x[q], twB[(q - 1) * 16 + t]);
    } else {
        x[0] = blk[t];
#pragma unroll
        for (int q = 1; q < 16; ++q) x[q] = cmulc(blk[16 * q + (t ^ q)], twB[(q - 1) * 16 + t]);
        CDft<16, true>::run(x);
#pragma unroll
        for (int j = 0; j < 16; ++j) blk[16 * j + (t ^ j)] = x[j];
    }
}

// ---- mid: item u in [0, 16*R0).  kt[q * (16*R0) + u] = Re K[k0 + R0*(k1 + 16 q)] / L with
// u = 16*k0 + k1.
template <int R0>
BLUR_HD void ols_mid(float2* __restrict__ s, int u, const float* __restrict__ kt)
{
    const int k1 = u & 15;
    float2* grp = s + 16 * u;
    float2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = grp[j ^ k1];
    CDft<16, false>::run(x);
#pragma unroll
    for (int q = 0; q < 16; ++q) x[q] = cscale(x[q], kt[q * (16 * R0) + u]);
    CDft<16, true>::run(x);
#pragma unroll
    for (int j = 0; j < 16; ++j) grp[j ^ k1] = x[j];
}

// ---- row pass IO: uint8 interleaved BGR lines -> Q -------------------------------------------
//
// HBM layouts of the overlap-save path (reference: pffft_, Source.cpp:429-570):
//   in   uint8 [rows][cols][3]
//   Q    float [3][nparts][rqs][cpp][4][2]    row-pass result = the reference's `resf` (:536) and
//        its flip_block (:540) in one: [.. rq][cp] is a 32-byte sector holding rows 4rq..4rq+3 of
//        columns 2cp, 2cp+1.  Row-pass CTAs (two rows each) write half sectors that merge in L2;
//        the column pass pulls a column pair's strip with TMA boxes of {8 floats, 128 sectors},
//        which land in shared memory as a contiguous (col 2cp, col 2cp+1) float2 sequence over
//        rows — exactly the packed input of its FFT.  `part` = destination GPU of the column pair
//        (1 on a single GPU), rqs = row quads per stripe, cpp = column pairs per part.
//   U    uint8 [3][ucols][rows]              column-pass result, planar transposed
//
// unit = (block * 3 + ch) * npairs + pair ; lines (2*pair, 2*pair + 1) of the stripe.
struct OlsRowIO {
    static constexpr bool kSmemInput = false;
    const uint8_t* in;     // stripe base (local row 0)
    float* q;              // Q base of this stripe
    int rows;              // valid rows in this stripe
    int cols;
    int npairs;            // ceil(rows / 2)
    int rqs;               // row quads per stripe (Q capacity)
    int cpp;               // column pairs per part
    int nparts;
    int nunits;
    OlsAxis ax;            // len = cols

    struct Ctx {
        const uint8_t* pa;   // line a, channel ch, pixel 0
        float* qa;           // Q float of (ch, part 0, rq, cp 0, row a & 3, col 0)
        int s;               // extended-line position of tile index 0
        int nvalid;          // outputs this block keeps: tile indices [halo, halo + nvalid)
        bool has_b, interior;
    };
    BLUR_HD Ctx begin(int unit) const
    {
        Ctx c;
        const int pair = unit % npairs, rest = unit / npairs;
        const int ch = rest % 3, b = rest / 3;
        const int ra = 2 * pair;
        c.has_b = ra + 1 < rows;
        c.pa = in + static_cast<size_t>(ra) * cols * 3 + ch;
        c.qa = q + ((static_cast<size_t>(ch) * nparts * rqs + (ra >> 2)) * cpp) * 8 + (ra & 3) * 2;
        c.s = b * ax.V - ax.halo;
        c.nvalid = cols - b * ax.V < ax.V ? cols - b * ax.V : ax.V;
        c.interior = c.s >= 0 && c.s + ax.L <= cols;
        return c;
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
    template <int R0>
    BLUR_HD void load_item(const Ctx& c, int t, float2 (&x)[R0]) const
    {
        const size_t rowb = static_cast<size_t>(cols) * 3;
        if (c.interior) {
            const uint8_t* p = c.pa + (c.s + t) * 3;
            if (c.has_b) {
#pragma unroll
                for (int j = 0; j < R0; ++j) x[j] = make_float2(u8f(p[768 * j]), u8f(p[rowb + 768 * j]));
            } else {
#pragma unroll
                for (int j = 0; j < R0; ++j) x[j] = make_float2(u8f(p[768 * j]), 0.f);
            }
        } else {
#pragma unroll
            for (int j = 0; j < R0; ++j) {
                const int g = ols_src(c.s + 256 * j + t, cols, ax.pad);
                x[j] = make_float2(0.f, 0.f);
                if (g >= 0) {
                    const uint8_t* p = c.pa + g * 3;
                    x[j] = make_float2(u8f(p[0]), c.has_b ? u8f(p[rowb]) : 0.f);
                }
            }
        }
    }
    template <int R0>
    BLUR_HD void store_item(const Ctx& c, int t, const float2 (&x)[R0]) const
    {
        int jlo, jhi;
        ols_item_range(t, ax.halo, c.nvalid, jlo, jhi);
        const int x0 = c.s + t;
        if (nparts == 1) {
            float* d = c.qa + static_cast<size_t>(x0 >> 1) * 8 + (x0 & 1);
#pragma unroll
            for (int j = 0; j < R0; ++j)
                if (j >= jlo && j <= jhi) {
                    d[1024 * j] = x[j].x;
                    if (c.has_b) d[1024 * j + 2] = x[j].y;
                }
        } else {
#pragma unroll
            for (int j = 0; j < R0; ++j)
                if (j >= jlo && j <= jhi) {
                    const int cp = (x0 >> 1) + 128 * j;
                    const int part = cp / cpp;
                    float* d = c.qa + (static_cast<size_t>(part) * rqs * cpp + (cp - part * cpp)) * 8 + (x0 & 1);
                    d[0] = x[j].x;
                    if (c.has_b) d[2] = x[j].y;
                }
        }
    }
};

// ---- column pass IO: Q -> U ------------------------------------------------------------------
// unit = (block * 3 + ch) * ncp + cp.  The block's L rows of the column pair arrive in shared
// memory by TMA (kernel) or a plain copy (emulator); rows outside the image are then patched
// with their reflect-101 source (or zero) by ols_col_patch.
struct OlsColIO {
    static constexpr bool kSmemInput = true;
    uint8_t* u;            // U base [3][ucols][rows]
    int rows;              // image rows = line length
    int ncp;               // column pairs handled here (capacity)
    int valid_ncp;         // pairs that exist in the image
    int ucols;             // 2 * ncp
    int nunits;
    OlsAxis ax;            // len = rows

    struct Ctx {
        uint8_t* ua;       // U line of column 2cp
        int s;             // row of tile index 0 (multiple of 4, may be negative)
        int nvalid;        // outputs this block keeps: tile indices [halo, halo + nvalid)
        int ch, cp;
        bool valid, interior;
    };
    BLUR_HD Ctx begin(int unit) const
    {
        Ctx c;
        c.cp = unit % ncp;
        const int rest = unit / ncp;
        c.ch = rest % 3;
        const int b = rest / 3;
        c.valid = c.cp < valid_ncp;
        c.ua = u + (static_cast<size_t>(c.ch) * ucols + 2 * c.cp) * rows;
        c.s = b * ax.V - ax.halo;
        c.nvalid = rows - b * ax.V < ax.V ? rows - b * ax.V : ax.V;
        c.interior = c.s >= 0 && c.s + ax.L <= rows;
        return c;
    }
    BLUR_HD static uint8_t to_u8(float v)
    {
        // (uint8_t)(v + 0.5f): truncation, no clamp (Utils.hpp:189,204-206); negatives (undefined
        // in the reference) saturate to 0
        const float t = v + 0.5f;
        return t > 0.f ? static_cast<uint8_t>(static_cast<int>(t)) : static_cast<uint8_t>(0);
    }
    template <int R0>
    BLUR_HD void load_item(const Ctx&, int, float2 (&)[R0]) const {}
    template <int R0>
    BLUR_HD void store_item(const Ctx& c, int t, const float2 (&x)[R0]) const
    {
        int jlo, jhi;
        ols_item_range(t, ax.halo, c.nvalid, jlo, jhi);
        uint8_t* d = c.ua + (c.s + t);
#pragma unroll
        for (int j = 0; j < R0; ++j)
            if (j >= jlo && j <= jhi) {
                d[256 * j] = to_u8(x[j].x);
                d[rows + 256 * j] = to_u8(x[j].y);
            }
    }
};

// Patch tile entry i of an edge block (natural order in shared memory): rows outside the image
// take their reflect-101 source, which is inside the same tile, or zero.
BLUR_HD void ols_col_patch(float2* __restrict__ s, int i, int s0, const OlsAxis& ax)
{
    const int g = s0 + i;
    if (g >= 0 && g < ax.len) return;
    const int src = ols_src(g, ax.len, ax.pad);
    s[i] = src < 0 ? make_float2(0.f, 0.f) : s[src - s0];
}

}  // namespace blurb200
