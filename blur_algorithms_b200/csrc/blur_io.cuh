// blur_io.cuh — global-memory front/back ends of the FFT convolution passes.
//
// Data layout in HBM for the Gaussian 1-D tile path (reference: pffft_, Source.cpp:429-570):
//
//   in   uint8  [rows][cols][3]                      interleaved BGR, as cv::Mat::data
//   Q    float  [part][3][cpp][qrows][2]             row-pass result.  Column pairs are the slow
//                                                    index, rows the fast one, so that the column
//                                                    pass reads each (channel, column-pair) strip
//                                                    as ONE contiguous run of float2 =
//                                                    (column 2cp, column 2cp+1) — i.e. exactly the
//                                                    packed complex input of its FFT.  `part`
//                                                    splits the column pairs by destination GPU
//                                                    (1 part on a single GPU); cpp = column pairs
//                                                    per part; qrows = row capacity of the stripe.
//                                                    This replaces resf + flip_block (:536,:540).
//   U    uint8  [3][ucols][rows]                     column-pass result after +0.5f truncation
//                                                    (Utils.hpp:204-206), planar and transposed;
//                                                    interleave_u8_kernel turns it into
//   out  uint8  [rows][cols][3]                      (replaces flip_block :562 + interleave :567)
//
// Tile placement: the reference builds  left-reflect | line | right-reflect | zeros
// (Source.cpp:525-529).  Circular convolution is shift invariant, so we rotate that tile left by
// `pad`:  line | right-reflect | zeros | left-reflect.  Output sample x is then at tile index x
// and all global accesses are naturally aligned.  (Result identical up to float rounding.)
#pragma once
#include "fft_core.cuh"

namespace blurb200 {

// tile index n -> source index in a line of length L (or -1 for the zero region)
BLUR_DEV_INLINE int tile_src(int n, int L, int pad, int N)
{
    if (n < L) return n;
    if (n < L + pad) return 2 * (L - 1) - n;     // right reflect-101
    if (n >= N - pad) return N - n;              // left reflect-101, wrapped to the tile's end
    return -1;                                   // trailing zeros
}

// ---- row pass: uint8 interleaved image lines -> Q --------------------------------------------
// IO functor contract used by fwd_pass<.., FIRST> / inv_pass<.., LAST>:
//   Ctx  begin(int unit) const;                              once per work item
//   void load (const Ctx&, int n, float& re, float& im);     n in [0, N)
//   void store(const Ctx&, int n, float re, float im);
//
// unit = (quad*3 + ch)*2 + pair ; row pair rp = 2*quad + pair ; lines (2rp, 2rp+1).
// Two consecutive units are the two row pairs that complete the same 32-byte sectors of Q.
struct RowIO {
    const uint8_t* in;     // stripe base: local row 0
    float* q;              // Q base of this stripe
    int rows;              // valid rows in this stripe
    int cols;
    int pad, N;
    int qrows;             // row capacity of Q (>= rows)
    int cpp;               // column pairs per part
    int nunits;

    struct Ctx {
        const uint8_t* pa;   // line a, channel ch, pixel 0
        float* qa;           // Q element (part 0, ch, cp 0, row a, col-in-pair 0)
        bool valid, has_b;
    };
    BLUR_DEV_INLINE Ctx begin(int unit) const
    {
        Ctx c;
        const int quad = unit / 6, rem = unit - quad * 6;
        const int ch = rem >> 1;
        const int ra = 2 * (2 * quad + (rem & 1));
        c.valid = ra < rows;
        c.has_b = ra + 1 < rows;
        c.pa = in + static_cast<size_t>(ra) * cols * 3 + ch;
        c.qa = q + (static_cast<size_t>(ch) * cpp * qrows + ra) * 2;
        return c;
    }
    BLUR_DEV_INLINE void load(const Ctx& c, int n, float& re, float& im) const
    {
        re = 0.f; im = 0.f;
        if (!c.valid) return;
        const int x = tile_src(n, cols, pad, N);
        if (x < 0) return;
        const uint8_t* p = c.pa + x * 3;
        re = static_cast<float>(p[0]);
        if (c.has_b) im = static_cast<float>(p[static_cast<size_t>(cols) * 3]);
    }
    BLUR_DEV_INLINE void store(const Ctx& c, int n, float re, float im) const
    {
        // an odd image width leaves a phantom column in the last pair: zero it so the packed
        // column FFT never mixes uninitialised memory into the real column
        if (n > cols || (n == cols && !(cols & 1)) || !c.valid) return;
        if (n == cols) { re = 0.f; im = 0.f; }
        const int cp = n >> 1;
        const int part = cp / cpp, cpl = cp - part * cpp;
        float* d = c.qa + (static_cast<size_t>(part) * 3 * cpp + cpl) * qrows * 2 + (n & 1);
        d[0] = re;
        if (c.has_b) d[2] = im;
    }
};

// ---- column pass: Q -> U ---------------------------------------------------------------------
// unit = ch * ncp + cp (cp local to this GPU's column-pair range).  Q is given as `nchunks`
// buffers (one per source stripe / GPU), each [3][ncp][chunk_rows][2]; chunk c holds rows
// [c*chunk_rows, ...).
struct ColIO {
    const float2* q;       // chunk 0 base; chunk c at q + c * chunk_stride
    size_t chunk_stride;   // in float2
    int chunk_rows;        // row capacity per chunk
    uint8_t* u;            // U base [3][ucols][rows]
    int rows;              // total rows (line length of the column pass)
    int ncp;               // column pairs handled here (capacity)
    int valid_ncp;         // pairs that exist in the image (<= ncp; the last rank may own fewer)
    int ucols;             // 2*ncp
    int pad, N;
    int nunits;

    struct Ctx {
        const float2* qs;    // strip base in chunk 0
        uint8_t* ua;         // U line of column 2cp
        bool valid;
    };
    BLUR_DEV_INLINE Ctx begin(int unit) const
    {
        Ctx c;
        const int ch = unit / ncp, cp = unit - ch * ncp;
        c.valid = unit < nunits && cp < valid_ncp;
        c.qs = q + static_cast<size_t>(unit) * chunk_rows;
        c.ua = u + (static_cast<size_t>(ch) * ucols + 2 * cp) * rows;
        return c;
    }
    BLUR_DEV_INLINE void load(const Ctx& c, int n, float& re, float& im) const
    {
        re = 0.f; im = 0.f;
        if (!c.valid) return;
        const int r = tile_src(n, rows, pad, N);
        if (r < 0) return;
        const int k = r / chunk_rows, rl = r - k * chunk_rows;
        const float2 v = c.qs[k * chunk_stride + rl];
        re = v.x; im = v.y;
    }
    BLUR_DEV_INLINE static uint8_t to_u8(float v)
    {
        // (uint8_t)(v + 0.5f): truncation, no clamp (Utils.hpp:189,204-206).  Negative inputs are
        // undefined in the reference; we saturate them to 0 like its DFT_image fixtures show.
        const float t = v + 0.5f;
        return t > 0.f ? static_cast<uint8_t>(static_cast<int>(t)) : static_cast<uint8_t>(0);
    }
    BLUR_DEV_INLINE void store(const Ctx& c, int n, float re, float im) const
    {
        if (n >= rows || !c.valid) return;
        c.ua[n] = to_u8(re);
        c.ua[rows + n] = to_u8(im);
    }
};

}  // namespace blurb200
