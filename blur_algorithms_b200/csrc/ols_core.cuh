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
// Shared-memory layout: F transforms (F = 1, 2 or 4) are interleaved element-wise and logical
// element p of transform X lives at slot F*(p ^ ((p >> 4) & M)) + X with M = 16/F - 1.  The XOR
// only touches the low bits, so
//   pass A : slot = F*(256*q + (t ^ (t>>4 & M))) + X        one XOR per item, immediates after that
//   pass B : slot = F*(256*k0 + 16*j + (t ^ (j & M))) + X
//   mid    : slot = F*(16*u + (j ^ (u & M))) + X
// Work item i of a pass is (X = i % F, index = i / F): the 16 lanes of a half-warp then cover 16/F
// consecutive indices x F transforms = 16 distinct 8-byte bank pairs in every pass: conflict
// free without padding.  A block that arrives by TMA in NATURAL order F*p + X (column pass: F
// column pairs = 2F adjacent floats per image row) is consumed by pass A in place: item (X, t)
// reads slots F*(256*j + t) + X and writes F*(256*q + swz(t)) + X; swz(t) stays inside t's
// aligned group of 16/F indices, i.e. inside the same half-warp of the same round, whose loads all
// precede its stores.
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

template <int F> BLUR_HD int ols_swz(int t) { return t ^ ((t >> 4) & (16 / F - 1)); }

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
//   void load_item <R0>(const Ctx&, int X, int t, float2 (&x)[R0]) const    x[j] = tile_X[256 j + t]
//   void store_item<R0>(const Ctx&, int X, int t, const float2 (&x)[R0]) const
// The valid outputs of a block are tile indices [halo, halo + nvalid); item t owns j in
// [jlo, jhi] of them (ols_item_range), so the per-element test is a compare of two registers
// with an immediate.
// In all pass functions `sx` = shared-memory base of the CTA's tile + X.

// ---- pass A, forward: item (X, t), t in [0, 256)
// `nat` = where a kSmemInput block sits in natural order: the kernel passes sx itself (in place,
// see the layout note above); the sequential CPU emulator passes a copy.
template <int R0, int F, class IO>
BLUR_HD void ols_pass_a_fwd(float2* sx, const float2* nat, int X, int t, const float2* __restrict__ twA, const IO& io,
                            const typename IO::Ctx& ctx)
{
    float2 x[R0];
    if (IO::kSmemInput) {
#pragma unroll
        for (int j = 0; j < R0; ++j) x[j] = nat[F * (256 * j + t)];
    } else {
        io.template load_item<R0>(ctx, X, t, x);
    }
    CDft<R0, false>::run(x);
    const int ta = ols_swz<F>(t);
    sx[F * ta] = x[0];
#pragma unroll
    for (int q = 1; q < R0; ++q) sx[F * (256 * q + ta)] = cmul(x[q], twA[(q - 1) * 256 + t]);
}

// ---- pass A, backward
template <int R0, int F, class IO>
BLUR_HD void ols_pass_a_inv(const float2* __restrict__ sx, int X, int t, const float2* __restrict__ twA, const IO& io,
                            const typename IO::Ctx& ctx)
{
    float2 x[R0];
    const int ta = ols_swz<F>(t);
    x[0] = sx[F * ta];
#pragma unroll
    for (int q = 1; q < R0; ++q) x[q] = cmulc(sx[F * (256 * q + ta)], twA[(q - 1) * 256 + t]);
    CDft<R0, true>::run(x);
    io.template store_item<R0>(ctx, X, t, x);
}

// ---- pass B: item (X, it), it in [0, 16*R0): k0 = it >> 4, t = it & 15
template <bool INV, int F>
BLUR_HD void ols_pass_b(float2* __restrict__ sx, int it, const float2* __restrict__ twB)
{
    constexpr int M = 16 / F - 1;
    const int t = it & 15;
    float2* blk = sx + F * ((it >> 4) << 8);
    float2 x[16];
    if (!INV) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = blk[F * (16 * j + (t ^ (j & M)))];
        CDft<16, false>::run(x);
        blk[F * t] = x[0];
#pragma unroll
        for (int q = 1; q < 16; ++q) blk[F * (16 * q + (t ^ (q & M)))] = cmul(x[q], twB[(q - 1) * 16 + t]);
    } else {
        x[0] = blk[F * t];
#pragma unroll
        for (int q = 1; q < 16; ++q) x[q] = cmulc(blk[F * (16 * q + (t ^ (q & M)))], twB[(q - 1) * 16 + t]);
        CDft<16, true>::run(x);
#pragma unroll
        for (int j = 0; j < 16; ++j) blk[F * (16 * j + (t ^ (j & M)))] = x[j];
    }
}

// ---- mid: item (X, u), u in [0, 16*R0).  kt[q * (16*R0) + u] = Re K[k0 + R0*(k1 + 16 q)] / L
// with u = 16*k0 + k1.
template <int R0, int F>
BLUR_HD void ols_mid(float2* __restrict__ sx, int u, const float* __restrict__ kt)
{
    constexpr int M = 16 / F - 1;
    const int k1 = u & M;
    float2* grp = sx + F * 16 * u;
    float2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = grp[F * (j ^ k1)];
    CDft<16, false>::run(x);
#pragma unroll
    for (int q = 0; q < 16; ++q) x[q] = cscale(x[q], kt[q * (16 * R0) + u]);
    CDft<16, true>::run(x);
#pragma unroll
    for (int j = 0; j < 16; ++j) grp[F * (j ^ k1)] = x[j];
}

// ---- row pass IO: uint8 interleaved BGR lines -> Q -------------------------------------------
//
// HBM layouts of the overlap-save path (reference: pffft_, Source.cpp:429-570):
//   in   uint8 [rows][cols][3]
//   Q    float [3][nparts][srows][qpitch]     row-pass result, planar row-major = the reference's
//        `resf` (Source.cpp:536).  Row-pass warps write 128 contiguous bytes per store.  The
//        reference's flip_block (:540) is replaced by TMA: the column pass pulls, per CTA, an
//        {8 columns x 256 rows} box, which lands in shared memory as four column-pair sequences
//        interleaved element-wise — the layout its FFT passes work in.  `part` = destination GPU
//        of the column (1 on a single GPU), srows = rows per stripe, qpitch = columns per part
//        rounded up to 8.
//   U    uint8 [3][noct][rows][8]             column-pass result: per channel and column octet,
//        8 bytes per image row (rounded with +0.5f truncation, Utils.hpp:204-206)
//   out  uint8 [rows][cols][3]                interleave_oct_kernel: U -> out (replaces
//        flip_block :562 + interleave_BGR :567)
//
// unit = (block * 3 + ch) * npairs + pair ; lines (2*pair, 2*pair + 1) of the stripe.
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
    int nunits;
    OlsAxis ax;            // len = cols

    struct Ctx {
        const uint8_t* pa;   // line a, channel ch, pixel 0
        float* qa;           // Q float of (ch, part 0, row a, col 0)
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
        c.qa = q + (static_cast<size_t>(ch) * nparts * srows + ra) * qpitch;
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
    BLUR_HD void load_item(const Ctx& c, int, int t, float2 (&x)[R0]) const
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
    BLUR_HD void store_item(const Ctx& c, int, int t, const float2 (&x)[R0]) const
    {
        int jlo, jhi;
        ols_item_range(t, ax.halo, c.nvalid, jlo, jhi);
        const int x0 = c.s + t;
        if (nparts == 1) {
            float* d = c.qa + x0;
#pragma unroll
            for (int j = 0; j < R0; ++j)
                if (j >= jlo && j <= jhi) {
                    d[256 * j] = x[j].x;
                    if (c.has_b) d[qpitch + 256 * j] = x[j].y;
                }
        } else {
#pragma unroll
            for (int j = 0; j < R0; ++j)
                if (j >= jlo && j <= jhi) {
                    const int xx = x0 + 256 * j;
                    const int part = xx / cpart;
                    float* d = c.qa + static_cast<size_t>(part) * srows * qpitch + (xx - part * cpart);
                    d[0] = x[j].x;
                    if (c.has_b) d[qpitch] = x[j].y;
                }
        }
    }
};

// ---- column pass IO: Q -> U ------------------------------------------------------------------
// unit = (block * 3 + ch) * noct + oct.  One CTA transforms the four column pairs of a column
// octet at once (kF = 4).  The block's L rows x 8 columns arrive in shared memory by TMA (kernel)
// or a plain copy (emulator); rows outside the image are then patched with their reflect-101
// source (or zero) by ols_col_patch.
struct OlsColIO {
    static constexpr bool kSmemInput = true;
    static constexpr int kF = 4;
    uint8_t* u;            // U base [3][noct][rows][8]
    int rows;              // image rows = line length
    int noct;              // column octets handled here
    int nunits;
    OlsAxis ax;            // len = rows

    struct Ctx {
        uint8_t* ua;       // U byte of (ch, oct, row 0, 0)
        int s;             // row of tile index 0 (may be negative)
        int nvalid;        // outputs this block keeps: tile indices [halo, halo + nvalid)
        int ch, oct;
        bool interior;
    };
    BLUR_HD Ctx begin(int unit) const
    {
        Ctx c;
        c.oct = unit % noct;
        const int rest = unit / noct;
        c.ch = rest % 3;
        const int b = rest / 3;
        c.ua = u + (static_cast<size_t>(c.ch) * noct + c.oct) * rows * 8;
        c.s = b * ax.V - ax.halo;
        c.nvalid = rows - b * ax.V < ax.V ? rows - b * ax.V : ax.V;
        c.interior = c.s >= 0 && c.s + ax.L <= rows;
        return c;
    }
    BLUR_HD static uint32_t to_u8(float v)
    {
        // (uint8_t)(v + 0.5f): truncation, no clamp (Utils.hpp:189,204-206); negatives (undefined
        // in the reference) saturate to 0
        const float t = v + 0.5f;
        return t > 0.f ? static_cast<uint32_t>(static_cast<int>(t)) & 255u : 0u;
    }
    template <int R0>
    BLUR_HD void load_item(const Ctx&, int, int, float2 (&)[R0]) const {}
    template <int R0>
    BLUR_HD void store_item(const Ctx& c, int X, int t, const float2 (&x)[R0]) const
    {
        int jlo, jhi;
        ols_item_range(t, ax.halo, c.nvalid, jlo, jhi);
        uint16_t* d = reinterpret_cast<uint16_t*>(c.ua + (static_cast<ptrdiff_t>(c.s + t) * 8 + 2 * X));
#pragma unroll
        for (int j = 0; j < R0; ++j)
            if (j >= jlo && j <= jhi) d[1024 * j] = static_cast<uint16_t>(to_u8(x[j].x) | (to_u8(x[j].y) << 8));
    }
};

// Patch entry e = 4*i + X of an edge block (natural order in shared memory): rows outside the
// image take their reflect-101 source, which is inside the same tile, or zero.
BLUR_HD void ols_col_patch(float2* __restrict__ s, int e, int s0, const OlsAxis& ax)
{
    const int i = e >> 2, X = e & 3;
    const int g = s0 + i;
    if (g >= 0 && g < ax.len) return;
    const int src = ols_src(g, ax.len, ax.pad);
    s[e] = src < 0 ? make_float2(0.f, 0.f) : s[4 * (src - s0) + X];
}

}  // namespace blurb200
