// blur_b200.cu — C ABI (include/blur_b200.h), plan management and the auxiliary kernels.
//
// Hot-path kernels: fft_conv_kernel<RowIO>/<ColIO> (fft_kernels.cuh) plus interleave_u8_kernel,
// boxblur_h_kernel / boxblur_v_kernel below.  No cuFFT, no library calls on the measured path,
// no CPU fallback: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/blur_b200.h"
#include "fft_launch.hpp"
#include "fft_plan.hpp"
#include "ols_launch.hpp"
#include "ols_plan.hpp"
#include "direct_conv.hpp"

namespace blurb200 {
__attribute__((visibility("hidden"))) void note_launches(uint64_t n);   // counter of this library's own kernel launches (other TUs)
// boxblur.cu
cudaError_t launch_box_fused_pass(const uint8_t* in, uint8_t* out, int w, int h, int C, int r, int batch, size_t img_stride,
                                  cudaStream_t st);
}
using namespace blurb200;

namespace {

std::atomic<uint64_t> g_launches{0};
thread_local std::string g_cuda_err;

#define CU_TRY(expr)                                                                           \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            g_cuda_err = std::string(#expr) + ": " + cudaGetErrorString(_e);                   \
            return BLUR_B200_ECUDA;                                                            \
        }                                                                                      \
    } while (0)

inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ------------------------------------------------------------------ auxiliary kernels

// U [3][ucols][rows] (uint8, planar, transposed)  ->  out [rows][out_cols][3] (interleaved).
// Tile: 32 columns x 128 rows x 3 channels through shared memory.  Replaces the reference's
// second flip_block (Source.cpp:562) and interleave_BGR (:567) for data that is already uint8.
constexpr int kIlTileC = 32, kIlTileR = 128;
__global__ void __launch_bounds__(256)
interleave_u8_kernel(const uint8_t* __restrict__ u, uint8_t* __restrict__ out, int rows, int ucols,
                     int ncols /* valid columns to emit */, int out_cols /* row pitch of out, in pixels */)
{
    __shared__ uint8_t tile[3][kIlTileC][kIlTileR + 4];
    const int c0 = blockIdx.x * kIlTileC, r0 = blockIdx.y * kIlTileR;
    const int tid = threadIdx.x;
    const bool vec = (rows & 3) == 0;
    // load: each warp takes one (channel, column) line of 128 rows; lane reads 4 rows
    for (int line = tid >> 5; line < 3 * kIlTileC; line += 8) {
        const int ch = line / kIlTileC, c = line - ch * kIlTileC;
        const int lane = tid & 31;
        const int col = c0 + c;
        if (col >= ncols) continue;
        const uint8_t* src = u + (static_cast<size_t>(ch) * ucols + col) * rows + r0;
        if (vec && r0 + kIlTileR <= rows) {
            const uchar4 v = *reinterpret_cast<const uchar4*>(src + 4 * lane);
            *reinterpret_cast<uchar4*>(&tile[ch][c][4 * lane]) = v;
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int r = 4 * lane + k;
                if (r0 + r < rows) tile[ch][c][r] = src[r];
            }
        }
    }
    __syncthreads();
    // store: each row of the tile is (up to) 96 contiguous bytes
    const int nc = min(kIlTileC, ncols - c0);
    const int nbytes = nc * 3;
    for (int idx = tid; idx < kIlTileR * 96; idx += 256) {
        const int r = idx / 96, b = idx - r * 96;
        if (b >= nbytes || r0 + r >= rows) continue;
        const int c = b / 3, ch = b - c * 3;
        out[(static_cast<size_t>(r0 + r) * out_cols + c0) * 3 + b] = tile[ch][c][r];
    }
}

// Q [part][3][cpp][qrows][2] -> planar float [3][rows][cols]   (parity hook only)
__global__ void q_to_planar_kernel(const float* __restrict__ q, float* __restrict__ planar, int rows, int cols,
                                   int cpp, int qrows)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y, ch = blockIdx.z;
    if (x >= cols) return;
    const int cp = x >> 1, part = cp / cpp, cpl = cp - part * cpp;
    planar[(static_cast<size_t>(ch) * rows + r) * cols + x] =
        q[((static_cast<size_t>(part * 3 + ch) * cpp + cpl) * qrows + r) * 2 + (x & 1)];
}

// overlap-save U [3][noct][rows][8] (uint8) -> out [rows][out_cols][3].  Replaces the reference's
// second flip_block (Source.cpp:562) and interleave_BGR (:567).  No shared memory: a thread owns
// 4 image rows x 16 columns (two octets).  Per channel and octet its four rows are one 32-byte
// sector of U (two 128-bit loads); the 48 output bytes of each row are assembled with byte
// permutes and leave as three 128-bit stores, the lanes of a warp covering 1536 contiguous bytes.
__device__ __forceinline__ void il_octet(uint32_t b, uint32_t g, uint32_t r, uint32_t& w0, uint32_t& w1, uint32_t& w2)
{
    w0 = __byte_perm(__byte_perm(b, g, 0x1040), r, 0x3410);     // b0 g0 r0 b1
    w1 = __byte_perm(__byte_perm(g, b, 0x2601), r, 0x3250);     // g1 r1 b2 g2
    w2 = __byte_perm(__byte_perm(r, b, 0x3072), g, 0x3710);     // r2 b3 g3 r3
}

__global__ void __launch_bounds__(256)
interleave_oct_kernel(const uint8_t* __restrict__ u, uint8_t* __restrict__ out, int rows, int noct,
                      int ncols /* valid columns to emit */, int out_cols /* row pitch of out, in pixels */, int vec,
                      int row_begin /* multiple of 4 */, int row_end)
{
    const int op = blockIdx.x * 32 + threadIdx.x;          // octet pair
    const int r0 = row_begin + (blockIdx.y * 8 + threadIdx.y) * 4;     // first of 4 rows
    const int c0 = op * 16;
    if (c0 >= ncols || r0 >= row_end) return;
    if (vec && c0 + 16 <= ncols && r0 + 4 <= row_end) {
        uint4 v[3][2][2];                                   // [channel][octet][row pair]: x,y = row 2k, z,w = row 2k+1
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const uint4* p = reinterpret_cast<const uint4*>(u + ((static_cast<size_t>(ch) * noct + 2 * op + e) * rows + r0) * 8);
                v[ch][e][0] = __ldg(p);
                v[ch][e][1] = __ldg(p + 1);
            }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t w[12];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const uint4 B = v[0][e][k >> 1], G = v[1][e][k >> 1], R = v[2][e][k >> 1];
                const uint32_t blo = (k & 1) ? B.z : B.x, bhi = (k & 1) ? B.w : B.y;
                const uint32_t glo = (k & 1) ? G.z : G.x, ghi = (k & 1) ? G.w : G.y;
                const uint32_t rlo = (k & 1) ? R.z : R.x, rhi = (k & 1) ? R.w : R.y;
                il_octet(blo, glo, rlo, w[6 * e + 0], w[6 * e + 1], w[6 * e + 2]);
                il_octet(bhi, ghi, rhi, w[6 * e + 3], w[6 * e + 4], w[6 * e + 5]);
            }
            uint4* d = reinterpret_cast<uint4*>(out + (static_cast<size_t>(r0 + k) * out_cols + c0) * 3);
            d[0] = make_uint4(w[0], w[1], w[2], w[3]);
            d[1] = make_uint4(w[4], w[5], w[6], w[7]);
            d[2] = make_uint4(w[8], w[9], w[10], w[11]);
        }
        return;
    }
    // ragged edge / unaligned pitch: byte by byte
    for (int k = 0; k < 4 && r0 + k < row_end; ++k)
        for (int e = 0; e < 2 && 2 * op + e < noct; ++e)
            for (int px = 0; px < 8; ++px) {
                const int c = c0 + 8 * e + px;
                if (c >= ncols) break;
                for (int ch = 0; ch < 3; ++ch)
                    out[(static_cast<size_t>(r0 + k) * out_cols + c) * 3 + ch] =
                        u[((static_cast<size_t>(ch) * noct + 2 * op + e) * rows + r0 + k) * 8 + px];
            }
}

// overlap-save Q [3][nparts][srows][qpitch] -> planar float [3][rows][cols]   (parity hook only)
__global__ void q_ols_to_planar_kernel(const float* __restrict__ q, float* __restrict__ planar, int rows, int cols,
                                       int qpitch, int cpart, int srows, int nparts)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y, ch = blockIdx.z;
    if (x >= cols) return;
    const int part = nparts == 1 ? 0 : x / cpart, xl = x - part * cpart;
    planar[(static_cast<size_t>(ch) * rows + r) * cols + x] =
        q[((static_cast<size_t>(ch) * nparts + part) * srows + r) * qpitch + xl];
}

// Reflect_101<uint8_t,C> (Utils.hpp:212-243): one thread per output byte.
__global__ void reflect101_u8_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int rows, int cols,
                                     int C, int top, int bottom, int left, int right)
{
    const int prow = rows + top + bottom, pcolb = (cols + left + right) * C;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (b >= pcolb || i >= prow) return;
    const int px = b / C, c = b - px * C;
    int sr = i - top;
    if (sr < 0) sr = -sr;
    if (sr >= rows) sr = 2 * (rows - 1) - sr;
    int sx = px - left;
    if (sx < 0) sx = -sx;
    if (sx >= cols) sx = 2 * (cols - 1) - sx;
    out[static_cast<size_t>(i) * pcolb + b] = in[(static_cast<size_t>(sr) * cols + sx) * C + c];
}

// deinterleave_BGR<uint8_t,float> (Utils.hpp:159-184)
__global__ void deinterleave_u8_f32_kernel(const uint8_t* __restrict__ in, float* __restrict__ b, float* __restrict__ g,
                                           float* __restrict__ r, uint32_t total)
{
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= total) return;
    b[x] = in[3 * static_cast<size_t>(x) + 0];
    g[x] = in[3 * static_cast<size_t>(x) + 1];
    r[x] = in[3 * static_cast<size_t>(x) + 2];
}

// interleave_BGR<uint8_t,float> (Utils.hpp:186-210): +0.5f, truncation
__global__ void interleave_f32_u8_kernel(const float* __restrict__ b, const float* __restrict__ g,
                                         const float* __restrict__ r, uint8_t* __restrict__ out, uint32_t total)
{
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= total) return;
    out[3 * static_cast<size_t>(x) + 0] = ColIO::to_u8(b[x]);
    out[3 * static_cast<size_t>(x) + 1] = ColIO::to_u8(g[x]);
    out[3 * static_cast<size_t>(x) + 2] = ColIO::to_u8(r[x]);
}

// flip_block<float,1>: in h x w -> out w x h, 32x32 tiles through padded shared memory
__global__ void transpose_f32_kernel(const float* __restrict__ in, float* __restrict__ out, int w, int h)
{
    __shared__ float tile[32][33];
    const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int x = x0 + threadIdx.x, y = y0 + j;
        if (x < w && y < h) tile[j][threadIdx.x] = in[static_cast<size_t>(y) * w + x];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int y = y0 + threadIdx.x, x = x0 + j;
        if (x < w && y < h) out[static_cast<size_t>(x) * h + y] = tile[threadIdx.x][j];
    }
}

// ---- pffft Nyquist quirk on the overlap-save path (Source.cpp:414-427) -----------
// The reference's pffft_ result per line is  exact[x] + c * (-1)^x * A,  c = (K[0] - K[N/2]) / N,
// A = sum over the reflect-padded tile of tile[m] (-1)^m = sum_x w_x (-1)^x line[x] (the pad parity
// cancels), w_x = 1 + [1 <= x <= pad] + [len-1-pad <= x <= len-2].  These two kernels produce c * A
// per line; the FFT kernels add it in their epilogues.
__device__ __forceinline__ int quirk_weight(int x, int len, int pad)
{
    return 1 + ((x >= 1 && x <= pad) ? 1 : 0) + ((x >= len - 1 - pad && x <= len - 2) ? 1 : 0);
}

// unsigned bytes of a  x  signed bytes of b, accumulated
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int acc)
{
    int r;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(acc));
    return r;
}

// rows: exact integer sums over the uint8 line.  grid = rows, 128 threads.  corr[ch * pitch + row].
// vec: cols % 4 == 0 and a 4-byte aligned image -> a thread takes 4 pixels as three 32-bit words
// and weights them with dp4a (signed selector bytes = weight * sign of each pixel).
__global__ void __launch_bounds__(128)
quirk_rows_kernel(const uint8_t* __restrict__ in, int cols, int pad, double c, float* __restrict__ corr, int pitch, int vec)
{
    __shared__ int part[4][3];
    const int row = blockIdx.x, tid = threadIdx.x;
    const uint8_t* line = in + static_cast<size_t>(row) * cols * 3;
    int acc[3] = {0, 0, 0};
    if (vec) {
        const uint32_t* lw = reinterpret_cast<const uint32_t*>(line);
        for (int g = tid; g < cols / 4; g += 128) {
            const uint32_t w0 = __ldg(lw + 3 * g), w1 = __ldg(lw + 3 * g + 1), w2 = __ldg(lw + 3 * g + 2);
            const int x = 4 * g;
            const uint32_t s0 = static_cast<uint8_t>(quirk_weight(x, cols, pad)), s1 = static_cast<uint8_t>(-quirk_weight(x + 1, cols, pad)),
                           s2 = static_cast<uint8_t>(quirk_weight(x + 2, cols, pad)), s3 = static_cast<uint8_t>(-quirk_weight(x + 3, cols, pad));
            // bytes: w0 = B0 G0 R0 B1, w1 = G1 R1 B2 G2, w2 = R2 B3 G3 R3
            acc[0] = dp4a_us(w0, static_cast<int>(s0 | (s1 << 24)), acc[0]);
            acc[0] = dp4a_us(w1, static_cast<int>(s2 << 16), acc[0]);
            acc[0] = dp4a_us(w2, static_cast<int>(s3 << 8), acc[0]);
            acc[1] = dp4a_us(w0, static_cast<int>(s0 << 8), acc[1]);
            acc[1] = dp4a_us(w1, static_cast<int>(s1 | (s2 << 24)), acc[1]);
            acc[1] = dp4a_us(w2, static_cast<int>(s3 << 16), acc[1]);
            acc[2] = dp4a_us(w0, static_cast<int>(s0 << 16), acc[2]);
            acc[2] = dp4a_us(w1, static_cast<int>(s1 << 8), acc[2]);
            acc[2] = dp4a_us(w2, static_cast<int>(s2 | (s3 << 24)), acc[2]);
        }
    } else {
        for (int x = tid; x < cols; x += 128) {
            const int w = quirk_weight(x, cols, pad) * ((x & 1) ? -1 : 1);
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) acc[ch] += w * line[3 * x + ch];
        }
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        int v = acc[ch];
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        if ((tid & 31) == 0) part[tid >> 5][ch] = v;
    }
    __syncthreads();
    if (tid < 3) corr[static_cast<size_t>(tid) * pitch + row] =
        static_cast<float>(c * static_cast<double>(part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid]));
}

// columns: A_col[ch][x] = sum_r w_r (-1)^r Q[ch][r][x] without sweeping Q.  The row pass is linear,
// Q[r][x] = rowblur(in[r])[x] + (-1)^x corr_row[r], so
//   A_col[x] = rowblur(S)[x] + (-1)^x * sum_r w_r (-1)^r corr_row[r],   S[x] = sum_r w_r (-1)^r in[r][x]
// S is an exact integer line (one sweep over the uint8 image, 3 B/px), its reflect-101 row blur a
// direct convolution with the taps in double (cols x ksize products).
constexpr int kQuirkSeg = 64;      // maximum; a launch uses nseg = clamp(rows / 128, 1, kQuirkSeg) row segments
// byte columns j of the interleaved image: partial[seg][j] = sum over the segment's rows of w_r (-1)^r in[r][j].
// grid (ceil(pitch/4/128), kQuirkSeg), 128 threads, one 32-bit word column per thread (vec) or one byte column.
__global__ void __launch_bounds__(128)
quirk_insum_partial(const uint8_t* __restrict__ in_local, int rows, int pitch /* cols*3 */, int pad, int* __restrict__ partial, int vec,
                    int row0 /* image row of in_local's first row */, int nlocal /* rows present */)
{
    // segments cut the LOCAL rows; weights and parities use the image row index (a sharded plan holds a stripe)
    const int seg_rows = (nlocal + gridDim.y - 1) / gridDim.y;
    const int r_begin = row0 + blockIdx.y * seg_rows, r_end = min(row0 + nlocal, r_begin + seg_rows);
    const uint8_t* in = in_local - static_cast<ptrdiff_t>(row0) * pitch;
    const int j = (blockIdx.x * 128 + threadIdx.x) * (vec ? 4 : 1);
    if (j >= pitch) return;
    int* out = partial + static_cast<size_t>(blockIdx.y) * pitch + j;
    if (vec) {
        int a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll 4
        for (int r = r_begin; r < r_end; ++r) {
            const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(in + static_cast<size_t>(r) * pitch + j));
            const int w = (quirk_weight(r, rows, pad) * ((r & 1) ? -1 : 1)) & 0xff;
            a0 = dp4a_us(v, w, a0); a1 = dp4a_us(v, w << 8, a1); a2 = dp4a_us(v, w << 16, a2); a3 = dp4a_us(v, w << 24, a3);
        }
        out[0] = a0; out[1] = a1; out[2] = a2; out[3] = a3;
    } else {
        int a0 = 0;
        for (int r = r_begin; r < r_end; ++r)
            a0 += quirk_weight(r, rows, pad) * ((r & 1) ? -1 : 1) * in[static_cast<size_t>(r) * pitch + j];
        out[0] = a0;
    }
}

// corr_col[ch * qpitch + x] = c_col * ( sum_k taps[|k|] S[refl101(x + k)][ch] + (-1)^x rsum[ch] ),  k = -pad..pad,
// rsum[ch] = sum_r w_r (-1)^r corr_row[ch][r], S = sum of the segment partials.  grid (ceil(cols/128), 3),
// 128 threads; every CTA reduces rsum itself (fixed order) and stages its 128 + 2 pad values of S
// in shared memory (dynamic, (128 + 2 pad) ints).
__global__ void __launch_bounds__(128)
quirk_col_corr(const int* __restrict__ partial, int cols, int rows, int pad, const double* __restrict__ taps /* [pad+1] */,
               const float* __restrict__ corr_row, double c, float* __restrict__ corr, int qpitch, int nseg,
               int xoff /* image column of corr[.][0]: a sharded plan corrects its own column part */, int xend)
{
    extern __shared__ int sS[];
    __shared__ double red[128];
    const int tid = threadIdx.x, ch = blockIdx.y, x0 = xoff + blockIdx.x * 128;
    const int pitch = cols * 3;
    const float* cr = corr_row + static_cast<size_t>(ch) * rows;
    double a = 0;
    for (int r = tid; r < rows; r += 128) a += static_cast<double>(cr[r]) * (quirk_weight(r, rows, pad) * ((r & 1) ? -1 : 1));
    red[tid] = a;
    for (int e = tid; e < 128 + 2 * pad; e += 128) {
        int xs = x0 - pad + e;
        if (xs < 0) xs = -xs;
        if (xs >= cols) xs = 2 * (cols - 1) - xs;
        int sv = 0;
        if (xs >= 0 && xs < cols) {
#pragma unroll 4
            for (int g = 0; g < nseg; ++g) sv += partial[static_cast<size_t>(g) * pitch + 3 * xs + ch];
        }
        sS[e] = sv;
    }
    __syncthreads();
    for (int d = 64; d; d >>= 1) {
        if (tid < d) red[tid] += red[tid + d];
        __syncthreads();
    }
    const int x = x0 + tid;
    if (x >= cols || x >= xend) return;
    double acc = taps[0] * static_cast<double>(sS[tid + pad]);
    for (int k = 1; k <= pad; ++k) acc += taps[k] * (static_cast<double>(sS[tid + pad - k]) + static_cast<double>(sS[tid + pad + k]));
    acc += ((x & 1) ? -1.0 : 1.0) * red[0];
    corr[static_cast<size_t>(ch) * qpitch + (x - xoff)] = static_cast<float>(c * acc);
}

// ------------------------------------------------------------------ box blur kernels
// The fused one-kernel-per-pass path lives in boxblur.cu; the two kernels below are the general
// fallback (any channel count, any window that fits a row in shared memory).
// Declared semantics: out = (2*sum + k) / (2k), sum over the reflect-101 window.

struct DivK {
    uint32_t k, d, magic;   // d = 2k ; magic = ceil(2^32 / d), exact while (511k)*(2k) < 2^32
    int use_magic;
};
__device__ __forceinline__ uint32_t box_round(uint32_t sum, const DivK& dk)
{
    const uint32_t n = 2u * sum + dk.k;
    return dk.use_magic ? __umulhi(n, dk.magic) : n / dk.d;
}

__device__ __forceinline__ int cdiv_dev(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ int refl101(int i, int L)
{
    if (i < 0) i = -i;
    if (i >= L) i = 2 * (L - 1) - i;
    return i;
}

// Horizontal pass: one CTA per image row.  The row, extended by r reflected pixels on each side,
// is staged in shared memory as uint32; a block-wide prefix sum per channel (warp shuffles inside
// each warp, one shared array across warps) turns every window sum into a difference of two
// prefix values.
__global__ void __launch_bounds__(256)
boxblur_h_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int w, int h, int C, int r, DivK dk,
                 size_t img_stride)
{
    extern __shared__ uint32_t pre[];                 // (w + 2r) * C prefix values
    __shared__ uint32_t warp_tot[8][4];
    const int y = blockIdx.x;
    const uint8_t* row = in + blockIdx.y * img_stride + static_cast<size_t>(y) * w * C;
    uint8_t* orow = out + blockIdx.y * img_stride + static_cast<size_t>(y) * w * C;
    const int E = w + 2 * r;                            // extended length in pixels
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    // 1. extended row into shared memory
    for (int e = tid; e < E * C; e += 256) {
        const int j = e / C, c = e - j * C;
        pre[e] = row[refl101(j - r, w) * C + c];
    }
    __syncthreads();

    // 2. per-channel inclusive prefix sum: thread t owns pixels [t*ppt, (t+1)*ppt)
    const int ppt = cdiv_dev(E, 256);
    const int j0 = min(tid * ppt, E), j1 = min(j0 + ppt, E);
    uint32_t acc[4] = {0, 0, 0, 0};
    for (int j = j0; j < j1; ++j)
        for (int c = 0; c < C; ++c) {
            acc[c] += pre[j * C + c];
            pre[j * C + c] = acc[c];
        }
    uint32_t off[4];
    for (int c = 0; c < C; ++c) {
        uint32_t v = acc[c];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, v, d);
            if (lane >= d) v += n;
        }
        if (lane == 31) warp_tot[wid][c] = v;
        off[c] = v - acc[c];                             // exclusive within the warp
    }
    __syncthreads();
    for (int c = 0; c < C; ++c)
        for (int k = 0; k < wid; ++k) off[c] += warp_tot[k][c];
    for (int j = j0; j < j1; ++j)
        for (int c = 0; c < C; ++c) pre[j * C + c] += off[c];
    __syncthreads();

    // 3. window sums: sum(x) = P[x + 2r] - P[x - 1]
    for (int e = tid; e < w * C; e += 256) {
        const int x = e / C, c = e - x * C;
        const uint32_t hi = pre[(x + 2 * r) * C + c];
        const uint32_t lo = x ? pre[(x - 1) * C + c] : 0u;
        orow[e] = static_cast<uint8_t>(box_round(hi - lo, dk));
    }
}

// Vertical pass: one thread per byte column (4 adjacent bytes per thread when the row pitch is a
// multiple of 4), sliding accumulator down a segment of `seg` rows.
template <int V>
__global__ void __launch_bounds__(128)
boxblur_v_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int pitch /* w*C bytes */, int h, int r,
                 DivK dk, int seg, size_t img_stride)
{
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) * V;
    if (b >= pitch) return;
    const int y0 = blockIdx.y * seg, y1 = min(y0 + seg, h);
    const uint8_t* src = in + blockIdx.z * img_stride + b;
    uint8_t* dst = out + blockIdx.z * img_stride + b;
    uint32_t acc[V];
#pragma unroll
    for (int v = 0; v < V; ++v) acc[v] = 0;
    auto ld = [&](int y, uint32_t* o) {
        const uint8_t* p = src + static_cast<size_t>(refl101(y, h)) * pitch;
        if (V == 4) {
            const uchar4 q = *reinterpret_cast<const uchar4*>(p);
            o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w;
        } else {
            o[0] = p[0];
        }
    };
    for (int i = -r; i <= r; ++i) {
        uint32_t t[V];
        ld(y0 + i, t);
#pragma unroll
        for (int v = 0; v < V; ++v) acc[v] += t[v];
    }
    for (int y = y0; y < y1; ++y) {
        uint8_t* p = dst + static_cast<size_t>(y) * pitch;
        if (V == 4) {
            uchar4 o;
            o.x = box_round(acc[0], dk); o.y = box_round(acc[1], dk);
            o.z = box_round(acc[2], dk); o.w = box_round(acc[3], dk);
            *reinterpret_cast<uchar4*>(p) = o;
        } else {
            p[0] = static_cast<uint8_t>(box_round(acc[0], dk));
        }
        uint32_t a[V], s[V];
        ld(y + r + 1, a);
        ld(y - r, s);
#pragma unroll
        for (int v = 0; v < V; ++v) acc[v] += a[v] - s[v];
    }
}

}  // namespace

namespace blurb200 {
void note_launches(uint64_t n) { g_launches += n; }
// other TUs (layout_typed.cu) report their CUDA errors through blur_b200_last_cuda_error too
__attribute__((visibility("hidden"))) void set_cuda_error(const char* what, cudaError_t e)
{
    g_cuda_err = std::string(what) + ": " + cudaGetErrorString(e);
}
}

// ------------------------------------------------------------------ plan

struct AxisTables {
    FftGeom g;
    LaunchShape shape;
    float2* d_tw = nullptr;
    float* d_kt = nullptr;
};

struct blur_b200_plan {
    int device = 0;
    int rows = 0, cols = 0;
    double sigma = 0;
    int mode = 0;
    GaussGeometry gg{};
    int rank = 0, nranks = 1;
    int stripe_rows = 0, cpp = 0, row0 = 0, valid_rows = 0, valid_ncp = 0;
    size_t block_floats = 0;          // floats per (stripe, part) block of Q
    AxisTables row, col;
    int algo = 0;                     // 0 = full-length FFT per line, 1 = overlap-save blocks
    OlsAxis ols_row{}, ols_col{};
    float2 *d_ols_twa_row = nullptr, *d_ols_twa_col = nullptr, *d_ols_twb = nullptr;
    float *d_ols_kt_row = nullptr, *d_ols_kt_col = nullptr;
    float2* d_ols2_tw = nullptr;      // two-factor kernels (ols2_kernels.cuh): twiddles and multipliers, shared by both axes
    float* d_ols2_kt = nullptr;
    int qpitch = 0, cpart = 0;        // overlap-save Q: floats per row / image columns per part
    float* d_q = nullptr;             // [nranks] blocks
    uint8_t* d_u = nullptr;           // [3][2*cpp][rows]
    int* d_cnt = nullptr;             // fused interleave of the two-factor column kernel: arrival counters per item
    bool col_fuse = false;            // BLUR_B200_COL_FUSE=1 at plan creation
    // narrow kernels on one GPU: the whole-image entries (run_device, run_host*) use the fused direct-convolution
    // kernel (direct_conv.cu) instead of the two FFT passes; the pass-wise entries keep the FFT kernels
    bool direct = false;
    float* d_dtaps = nullptr;         // [pad + 1] taps at offsets 0..pad
    // pffft Nyquist quirk served by the overlap-save kernels (single GPU): per-line corrections
    bool quirk_ols = false;
    double quirk_c_row = 0, quirk_c_col = 0;      // (K[0] - K[N/2]) / N of each axis
    float *d_corr_row = nullptr, *d_corr_col = nullptr;   // [3][rows], [3][qpitch]
    int* d_insum = nullptr;                               // [kQuirkSeg][cols*3] partial column sums of the input
    double* d_taps = nullptr;                             // centred taps [pad+1]
    // streamed host path (blur_b200_gauss_run_host_batch), all lazy: input / output images double
    // buffered, one stream per PCIe direction, events (image parity x chunk)
    uint8_t* d_sin[2] = {};
    uint8_t* d_sout[2] = {};
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    static constexpr int kHostChunks = 8;
    cudaEvent_t ev_h2d[kHostChunks] = {}, ev_out[kHostChunks + 1] = {};
    cudaEvent_t ev_rowdone[2] = {}, ev_d2hdone[2] = {};
    float* d_q_recv = nullptr;        // unused on a single GPU
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    static constexpr int kMaxChunks = 16;
    cudaEvent_t ev_chunk[kMaxChunks] = {};
    cudaStream_t ce_stream[7] = {};                // one per destination GPU: copies to different peers run on different engines
    cudaEvent_t ev_ce[7] = {};
    // staged exchange (blur_b200_gauss_row_pass_staged): row launch of stage k done / copy of stage k to destination d done
    static constexpr int kMaxStages = 8;
    cudaEvent_t ev_stage[kMaxStages] = {};
    cudaEvent_t ev_scopy[kMaxStages][7] = {};
    bool staged_store = false;                     // last staged row pass stored to the peers itself (no copy-engine events)
};

namespace {

int build_axis(AxisTables& ax, int N, double sigma, int ksize, int mode)
{
    std::vector<int> ovr;
    if (const char* e = std::getenv("BLUR_B200_PLAN")) {
        // "17280:18,30;4096:8,16" -> strided radices per FFT length
        std::string s(e);
        size_t pos = 0;
        while (pos < s.size()) {
            size_t semi = s.find(';', pos);
            std::string item = s.substr(pos, semi == std::string::npos ? std::string::npos : semi - pos);
            size_t colon = item.find(':');
            if (colon != std::string::npos && std::atoi(item.substr(0, colon).c_str()) == N) {
                std::string rs = item.substr(colon + 1);
                size_t p2 = 0;
                while (p2 < rs.size()) {
                    size_t comma = rs.find(',', p2);
                    ovr.push_back(std::atoi(rs.substr(p2, comma == std::string::npos ? std::string::npos : comma - p2).c_str()));
                    if (comma == std::string::npos) break;
                    p2 = comma + 1;
                }
            }
            if (semi == std::string::npos) break;
            pos = semi + 1;
        }
    }
    if (!make_geom(N, ax.g, ovr.empty() ? nullptr : &ovr)) return BLUR_B200_ESIZE;
    ax.shape = choose_launch(ax.g, kMaxThreads);
    if (ax.shape.smem_bytes > 227 * 1024) return BLUR_B200_ESIZE;
    std::vector<float2> tw;
    make_twiddles(N, tw);
    std::vector<double> K;
    kernel_spectrum(sigma, ksize, N, K);
    std::vector<float> kt;
    make_kt(ax.g, K, mode, kt);
    CU_TRY(cudaMalloc(&ax.d_tw, sizeof(float2) * tw.size()));
    CU_TRY(cudaMalloc(&ax.d_kt, sizeof(float) * kt.size()));
    CU_TRY(cudaMemcpy(ax.d_tw, tw.data(), sizeof(float2) * tw.size(), cudaMemcpyHostToDevice));
    CU_TRY(cudaMemcpy(ax.d_kt, kt.data(), sizeof(float) * kt.size(), cudaMemcpyHostToDevice));
    return BLUR_B200_OK;
}

template <class T>
int upload(T** dst, const std::vector<T>& v)
{
    CU_TRY(cudaMalloc(dst, sizeof(T) * v.size()));
    CU_TRY(cudaMemcpy(*dst, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice));
    return BLUR_B200_OK;
}

int build_ols_tables(blur_b200_plan* p, double sigma, int ksize)
{
    std::vector<float2> tw;
    std::vector<float> kt;
    if (p->ols_row.R2 || p->ols_col.R2) {
        make_ols2_tables(p->ols_row.R2 ? p->ols_row.R2 : p->ols_col.R2, sigma, ksize, tw, kt);
        int rc2 = upload(&p->d_ols2_tw, tw);
        if (rc2) return rc2;
        if ((rc2 = upload(&p->d_ols2_kt, kt))) return rc2;
        if (p->ols_row.R2 && p->ols_col.R2) return BLUR_B200_OK;
    }
    make_ols_twiddles_b(tw);
    int rc = upload(&p->d_ols_twb, tw);
    if (rc) return rc;
    make_ols_twiddles_a(p->ols_row.R0, tw);
    if ((rc = upload(&p->d_ols_twa_row, tw))) return rc;
    make_ols_kt(p->ols_row.R0, sigma, ksize, kt);
    if ((rc = upload(&p->d_ols_kt_row, kt))) return rc;
    make_ols_twiddles_a(p->ols_col.R0, tw);
    if ((rc = upload(&p->d_ols_twa_col, tw))) return rc;
    make_ols_kt(p->ols_col.R0, sigma, ksize, kt);
    return upload(&p->d_ols_kt_col, kt);
}

int plan_create_impl(blur_b200_plan* p, const GaussGeometry& gg, int rows, int cols, double sigma, int mode, int device, int rank,
                     int nranks);

int plan_create(blur_b200_plan** out, int rows, int cols, double sigma, int mode, int device, int rank, int nranks)
{
    // sigma < 0: FFT box blur, sigma = -kLen (fft_plan.hpp, "kernel kinds"); only blur_b200_boxfft_* pass it
    if (!out || rows < 1 || cols < 1 || !(sigma > 0 || (sigma <= -3 && box_klen(sigma) % 2 == 1)) || (mode != 0 && mode != 1) ||
        nranks < 1 || rank < 0 || rank >= nranks)
        return BLUR_B200_EINVAL;
    *out = nullptr;
    const GaussGeometry gg = gauss_geometry(rows, cols, sigma);
    // The reference reads out of bounds when pad > dim-1 in the smaller dimension
    // (README.md:33-38); deliberate tightening: refuse.
    if (gg.pad < 0 || gg.pad > rows - 1 || gg.pad > cols - 1) return BLUR_B200_EPAD;
    CU_TRY(cudaSetDevice(device));
    if (const char* e = std::getenv("BLUR_B200_L2_FETCH"))      // experiment knob: L2 fetch granularity in bytes (32 / 64 / 128)
        CU_TRY(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, static_cast<size_t>(std::atoi(e))));
    blur_b200_plan* p = new (std::nothrow) blur_b200_plan();
    if (!p) return BLUR_B200_ENOMEM;
    const int rc = plan_create_impl(p, gg, rows, cols, sigma, mode, device, rank, nranks);
    if (rc != BLUR_B200_OK) { blur_b200_gauss_plan_destroy(p); return rc; }      // nothing leaks on a CUDA failure half way
    *out = p;
    return BLUR_B200_OK;
}

int plan_create_impl(blur_b200_plan* p, const GaussGeometry& gg, int rows, int cols, double sigma, int mode, int device, int rank,
                     int nranks)
{
    p->device = device; p->rows = rows; p->cols = cols; p->sigma = sigma; p->mode = mode; p->gg = gg;
    p->rank = rank; p->nranks = nranks;
    const ShardGeom sg = shard_geometry(gg, mode, rank, nranks);
    p->algo = sg.algo; p->ols_row = sg.ols_row; p->ols_col = sg.ols_col;
    p->cpp = sg.cpp; p->qpitch = sg.qpitch; p->cpart = sg.cpart;
    p->stripe_rows = sg.stripe_rows; p->row0 = sg.row0; p->valid_rows = sg.valid_rows; p->valid_ncp = sg.valid_ncp;
    p->block_floats = sg.block_floats;
    int rc = BLUR_B200_OK;
    if (p->algo == 1) {
        rc = build_ols_tables(p, sigma, gg.ksize);
        if (rc == BLUR_B200_OK && mode == 1) {
            // K[0] = sum of the taps, K[N/2] = alternating sum (the taps are symmetric about index 0)
            std::vector<float> taps;
            for (int axis = 0; axis < 2; ++axis) {
                const int N = axis ? gg.n_col : gg.n_row;
                get_gaussian(taps, sigma, gg.ksize, N);
                double k0 = 0, kn = 0;
                for (int n = 0; n < N; ++n) { k0 += taps[n]; kn += (n & 1) ? -static_cast<double>(taps[n]) : taps[n]; }
                // float products like the reference: kernel_dft[0] * scaler (Source.cpp:423,506-507)
                const float scaler = 1.f / static_cast<float>(N);
                (axis ? p->quirk_c_col : p->quirk_c_row) =
                    static_cast<double>(static_cast<float>(k0) * scaler) - static_cast<double>(static_cast<float>(kn) * scaler);
            }
            p->quirk_ols = true;
            CU_TRY(cudaMalloc(&p->d_corr_row, sizeof(float) * 3 * rows));
            CU_TRY(cudaMalloc(&p->d_corr_col, sizeof(float) * 3 * p->qpitch));
            CU_TRY(cudaMalloc(&p->d_insum, sizeof(int) * kQuirkSeg * 3 * static_cast<size_t>(cols)));
            CU_TRY(cudaMemset(p->d_corr_col, 0, sizeof(float) * 3 * p->qpitch));
            get_gaussian(taps, sigma, gg.ksize, gg.n_row);          // rotated: centre tap at 0, offset k at index k
            std::vector<double> tc(static_cast<size_t>(gg.pad) + 1);
            for (int k = 0; k <= gg.pad; ++k) tc[k] = taps[k];
            CU_TRY(cudaMalloc(&p->d_taps, sizeof(double) * tc.size()));
            CU_TRY(cudaMemcpy(p->d_taps, tc.data(), sizeof(double) * tc.size(), cudaMemcpyHostToDevice));
        }
    } else {
        rc = build_axis(p->row, gg.n_row, sigma, gg.ksize, mode);
        if (rc == BLUR_B200_OK) rc = build_axis(p->col, gg.n_col, sigma, gg.ksize, mode);
    }
    if (rc != BLUR_B200_OK) return rc;
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    p->num_sms = prop.multiProcessorCount;
    CU_TRY(cudaMalloc(&p->d_q, sizeof(float) * p->block_floats * nranks));
    CU_TRY(cudaMemset(p->d_q, 0, sizeof(float) * p->block_floats * nranks));   // phantom column of an odd width stays 0
    CU_TRY(cudaMalloc(&p->d_u, static_cast<size_t>(3) * 2 * p->cpp * rows));
    if (const char* e = std::getenv("BLUR_B200_COL_FUSE")) p->col_fuse = std::atoi(e) != 0;
    {
        // direct path: measured faster than the block FFTs up to 41 taps (BLUR_B200_DIRECT=0 keeps the FFT kernels)
        const char* e = std::getenv("BLUR_B200_DIRECT");
        const bool want = !(e && std::atoi(e) == 0);
        if (want && nranks == 1 && p->algo == 1 && gg.pad >= 1 && gg.pad <= kDirectMaxPad) {
            std::vector<float> taps;
            get_gaussian(taps, sigma, gg.ksize, gg.n_row);          // rotated: offset k at index k
            taps.resize(static_cast<size_t>(gg.pad) + 1);
            CU_TRY(cudaMalloc(&p->d_dtaps, sizeof(float) * taps.size()));
            CU_TRY(cudaMemcpy(p->d_dtaps, taps.data(), sizeof(float) * taps.size(), cudaMemcpyHostToDevice));
            p->direct = true;
        }
    }
    if (p->col_fuse && p->algo == 1 && p->ols_col.R2) {
        const size_t n = static_cast<size_t>(p->qpitch / 32 + 2) * p->ols_col.nblocks;
        CU_TRY(cudaMalloc(&p->d_cnt, sizeof(int) * n));
        CU_TRY(cudaMemset(p->d_cnt, 0, sizeof(int) * n));
    }
    CU_TRY(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    return BLUR_B200_OK;
}

// row / column FFT kernels of an overlap-save plan: two-factor (ols2) or three-factor instantiation
cudaError_t launch_row_fft(blur_b200_plan* p, const OlsRowIO& io_in, int num_sms, cudaStream_t st)
{
    OlsRowIO io = io_in;
    // unit order of the two-factor row grid (OlsRowIO::pair_major); BLUR_B200_ROW_ORDER=0/1 forces it (A/B measurements)
    static const int order_env = std::getenv("BLUR_B200_ROW_ORDER") ? std::atoi(std::getenv("BLUR_B200_ROW_ORDER")) : -1;
    io.pair_major = order_env >= 0 ? order_env : (io.peer ? 1 : 0);
    if (io.ax.R2) return launch_ols2_row(io, p->d_ols2_tw, p->d_ols2_kt, num_sms, st);
    return launch_ols_row(io, p->d_ols_twa_row, p->d_ols_twb, p->d_ols_kt_row, num_sms, st);
}

// row pairs per wave of the persistent row grid (chunk granularity of the streamed / sharded paths)
int row_wave_pairs(const blur_b200_plan* p)
{
    const int units = p->ols_row.R2 ? ols2_row_units_per_sm(p->ols_row.R2) : ols_resident_ctas(p->ols_row.R0);
    return std::max(1, p->num_sms * units / std::max(1, p->ols_row.nblocks));
}

// pffft-quirk plans: column corrections from the input image and the row corrections (all rows of
// d_corr_row must have been produced on `st` before)
// row segments of the integer column sums: the same count on every rank of a sharded plan (the partials are
// summed element-wise across the ranks)
int quirk_nseg(const blur_b200_plan* p) { return std::max(1, std::min(kQuirkSeg, p->stripe_rows / 128)); }

// local half: per-row corrections of this plan's rows (into d_corr_row at their image row index) and the partial
// integer column sums over them.  A sharded plan's caller then sums d_corr_row and d_insum over the ranks.
int quirk_local(blur_b200_plan* p, const uint8_t* d_in_stripe, cudaStream_t st)
{
    if (p->nranks > 1) CU_TRY(cudaMemsetAsync(p->d_corr_row, 0, sizeof(float) * 3 * p->rows, st));
    const int pitch = p->cols * 3;
    const int vec = pitch % 4 == 0 && reinterpret_cast<uintptr_t>(d_in_stripe) % 4 == 0;
    const int nseg = quirk_nseg(p);
    if (p->valid_rows > 0) {
        quirk_rows_kernel<<<p->valid_rows, 128, 0, st>>>(d_in_stripe, p->cols, p->gg.pad, p->quirk_c_row, p->d_corr_row + p->row0, p->rows, vec);
        CU_TRY(cudaGetLastError());
        ++g_launches;
    }
    quirk_insum_partial<<<dim3(cdiv(vec ? pitch / 4 : pitch, 128), nseg), 128, 0, st>>>(d_in_stripe, p->rows, pitch, p->gg.pad, p->d_insum, vec,
                                                                                       p->row0, std::max(0, p->valid_rows));
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

// second half: column corrections of this plan's column part from the (summed) d_insum and d_corr_row
int quirk_finish(blur_b200_plan* p, cudaStream_t st)
{
    const int xoff = p->nranks > 1 ? p->rank * p->cpart : 0;
    const int xend = p->nranks > 1 ? std::min(p->cols, xoff + p->cpart) : p->cols;
    if (xend <= xoff) return BLUR_B200_OK;
    quirk_col_corr<<<dim3(cdiv(xend - xoff, 128), 3), 128, sizeof(int) * (128 + 2 * p->gg.pad), st>>>(
        p->d_insum, p->cols, p->rows, p->gg.pad, p->d_taps, p->d_corr_row, p->quirk_c_col, p->d_corr_col, p->qpitch, quirk_nseg(p), xoff, xend);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

// single-GPU plans: both halves back to back (all rows of d_corr_row must have been produced on `st` before)
int quirk_prepare_cols(blur_b200_plan* p, const uint8_t* d_in, cudaStream_t st)
{
    const int pitch = p->cols * 3;
    const int vec = pitch % 4 == 0 && reinterpret_cast<uintptr_t>(d_in) % 4 == 0;
    quirk_insum_partial<<<dim3(cdiv(vec ? pitch / 4 : pitch, 128), quirk_nseg(p)), 128, 0, st>>>(d_in, p->rows, pitch, p->gg.pad, p->d_insum, vec,
                                                                                                0, p->rows);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return quirk_finish(p, st);
}

int run_row_pass(blur_b200_plan* p, const uint8_t* d_in_stripe, float* d_q, cudaStream_t st, float* const* peers = nullptr)
{
    if (p->valid_rows <= 0) return BLUR_B200_OK;
    if (p->algo == 1) {
        OlsRowIO io{};
        io.in = d_in_stripe; io.q = d_q;
        io.rows = p->valid_rows; io.cols = p->cols;
        io.npairs = cdiv(p->valid_rows, 2);
        io.srows = p->stripe_rows; io.qpitch = p->qpitch; io.cpart = p->cpart; io.nparts = p->nranks;
        io.ax = p->ols_row;
        io.nunits = io.ax.nblocks * io.npairs;
        if (p->quirk_ols && p->nranks == 1) {
            quirk_rows_kernel<<<p->valid_rows, 128, 0, st>>>(d_in_stripe, p->cols, p->gg.pad, p->quirk_c_row, p->d_corr_row, p->rows,
                                                             p->cols % 4 == 0 && reinterpret_cast<uintptr_t>(d_in_stripe) % 4 == 0);
            CU_TRY(cudaGetLastError());
            ++g_launches;
            io.corr = p->d_corr_row; io.corr_pitch = p->rows;
            const int rc = quirk_prepare_cols(p, d_in_stripe, st);
            if (rc) return rc;
        } else if (p->quirk_ols) {
            // sharded: blur_b200_gauss_quirk_local has produced this stripe's rows (the kernel indexes stripe rows)
            io.corr = p->d_corr_row + p->row0; io.corr_pitch = p->rows;
        }
        if (peers && p->nranks > 1) {
            io.peer = 1; io.rank = p->rank;
            for (int k = 0; k < p->nranks; ++k) io.qpeer[k] = peers[k];
        }
        CU_TRY(launch_row_fft(p, io, p->num_sms, st));
        ++g_launches;
        return BLUR_B200_OK;
    }
    RowIO io{};
    io.in = d_in_stripe; io.q = d_q;
    io.rows = p->valid_rows; io.cols = p->cols; io.pad = p->gg.pad; io.N = p->row.g.N;
    io.qrows = p->stripe_rows; io.cpp = p->cpp;
    io.nunits = cdiv(p->valid_rows, 4) * 6;
    const LaunchShape& s = p->row.shape;
    const int grid = std::min(cdiv(io.nunits, s.batch), p->num_sms * s.ctas_per_sm);
    CU_TRY(launch_fft_row(p->row.g, io, p->row.d_tw, p->row.d_kt, s.batch, grid, s.threads, s.smem_bytes, st));
    ++g_launches;
    return BLUR_B200_OK;
}

// blk_begin/blk_end: overlap-save blocks (of V output rows) to transform; blk_end < 0 = all
// Fused interleave of the two-factor column kernel (OlsColIO::out).  OFF by default: measured at 16384^2 sigma = 50 it
// is slower than the separate pass (column + interleave 2.26 ms; fused 3.01 ms with 16-byte stores, 2.59 ms with
// 32-byte stores): L2 turns over in ~55 us at this kernel's traffic, so the U bytes of the triplet's other two
// channels are back in DRAM by the time the last CTA arrives (DRAM reads 3.39 -> 4.39 GB, profiles/r02l_*).
// BLUR_B200_COL_FUSE=1 (read when the plan is created) switches it on for measurements and for its parity test.
bool col_fuse_enabled(const blur_b200_plan* p) { return p->col_fuse && p->algo == 1 && p->ols_col.R2 && p->d_cnt; }

// d_out != nullptr and a two-factor column axis: the kernel also interleaves (writes d_out, pitch out_cols pixels,
// ncols valid columns) and *fused is set; otherwise the caller runs run_interleave afterwards.
int run_col_pass_to_u(blur_b200_plan* p, const float* d_q_chunks, cudaStream_t st, int blk_begin = 0, int blk_end = -1,
                      uint8_t* d_out = nullptr, int ncols = 0, int out_cols = 0, bool* fused = nullptr)
{
    if (fused) *fused = false;
    if (p->algo == 1) {
        OlsColIO io{};
        io.u = p->d_u;
        io.rows = p->rows;
        io.noct = std::min(p->qpitch / 8, cdiv(std::max(0, p->cols - p->rank * p->cpart), 8));
        io.ax = p->ols_col;
        if (blk_end < 0) blk_end = io.ax.nblocks;
        io.b0 = blk_begin;
        io.nbl = blk_end - blk_begin;
        io.nunits = io.ax.R2 ? ols2_col_npairs(io.noct, io.nbl) : io.nbl * 3 * io.noct;
        if (p->quirk_ols) {
            io.corr = p->d_corr_col; io.corr_pitch = p->qpitch;
        }
        if (d_out && col_fuse_enabled(p)) {
            io.out = d_out; io.out_cols = out_cols; io.ncols = std::min(ncols, io.noct * 8); io.cnt = p->d_cnt;
            if (fused) *fused = true;
        }
        if (io.ax.R2)
            CU_TRY(launch_ols2_col(io, d_q_chunks, p->qpitch, p->nranks * p->stripe_rows, p->d_ols2_tw, p->d_ols2_kt, p->num_sms, st));
        else
            CU_TRY(launch_ols_col(io, d_q_chunks, p->qpitch, p->nranks * p->stripe_rows, p->d_ols_twa_col, p->d_ols_twb,
                                  p->d_ols_kt_col, p->num_sms, st));
        ++g_launches;
        return BLUR_B200_OK;
    }
    ColIO io{};
    io.q = reinterpret_cast<const float2*>(d_q_chunks);
    io.chunk_stride = p->block_floats / 2;
    io.chunk_rows = p->stripe_rows;
    io.u = p->d_u;
    io.rows = p->rows; io.ncp = p->cpp; io.valid_ncp = p->valid_ncp; io.ucols = 2 * p->cpp;
    io.pad = p->gg.pad; io.N = p->col.g.N;
    io.nunits = 3 * p->cpp;
    const LaunchShape& s = p->col.shape;
    const int grid = std::min(cdiv(io.nunits, s.batch), p->num_sms * s.ctas_per_sm);
    CU_TRY(launch_fft_col(p->col.g, io, p->col.d_tw, p->col.d_kt, s.batch, grid, s.threads, s.smem_bytes, st));
    ++g_launches;
    return BLUR_B200_OK;
}

// row_begin (multiple of 4) / row_end: rows to emit, overlap-save plans only; row_end < 0 = all
int run_interleave(blur_b200_plan* p, uint8_t* d_out, int ncols, int out_cols, cudaStream_t st, int row_begin = 0,
                   int row_end = -1)
{
    if (ncols <= 0) return BLUR_B200_OK;
    if (row_end < 0) row_end = p->rows;
    if (row_end <= row_begin) return BLUR_B200_OK;
    if (p->algo == 1) {
        const int noct = std::min(p->qpitch / 8, cdiv(std::max(0, p->cols - p->rank * p->cpart), 8));
        const int vec = (static_cast<size_t>(out_cols) * 3 % 16 == 0) && (reinterpret_cast<uintptr_t>(d_out) % 16 == 0) &&
                        (reinterpret_cast<uintptr_t>(p->d_u) % 32 == 0) && (p->rows % 4 == 0);
        dim3 g(cdiv(cdiv(noct, 2), 32), cdiv(cdiv(row_end - row_begin, 4), 8));
        interleave_oct_kernel<<<g, dim3(32, 8), 0, st>>>(p->d_u, d_out, p->rows, noct, std::min(ncols, noct * 8), out_cols, vec,
                                                         row_begin, row_end);
        CU_TRY(cudaGetLastError());
        ++g_launches;
        return BLUR_B200_OK;
    }
    dim3 grid(cdiv(ncols, kIlTileC), cdiv(p->rows, kIlTileR));
    interleave_u8_kernel<<<grid, 256, 0, st>>>(p->d_u, d_out, p->rows, 2 * p->cpp, ncols, out_cols);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

}  // namespace

// ------------------------------------------------------------------ C ABI

extern "C" {

const char* blur_b200_strerror(int code)
{
    switch (code) {
    case BLUR_B200_OK: return "ok";
    case BLUR_B200_EINVAL: return "invalid argument";
    case BLUR_B200_EPAD: return "kernel half-width exceeds min(rows, cols) - 1";
    case BLUR_B200_ESIZE: return "FFT length not plannable on one SM (valid sizes up to 28160)";
    case BLUR_B200_ECUDA: return "CUDA error";
    case BLUR_B200_ENOMEM: return "out of memory";
    case BLUR_B200_ESCRATCH: return "image batch exceeds the built-in no-allocation work buffer (BLUR_B200_NOALLOC_SCRATCH_BYTES = 32 MiB)";
    default: return "unknown error";
    }
}

const char* blur_b200_last_cuda_error(void) { return g_cuda_err.c_str(); }

int blur_b200_is_valid_size(int n) { return is_valid_size(n); }
int blur_b200_nearest_transform_size(int n) { return nearest_transform_size(n); }
int blur_b200_gaussian_window(double sigma, int max_width) { return gaussian_window(sigma, max_width); }

int blur_b200_get_gaussian(float* kernel, double sigma, int width, int fft_length)
{
    if (!kernel || width < 0 || fft_length < 0) return BLUR_B200_EINVAL;
    const int eff_width = width ? width : gaussian_window(sigma);
    if (fft_length && fft_length < eff_width) return BLUR_B200_EINVAL;     // the reference would write past the vector (Source.cpp:80-89)
    std::vector<float> k;
    get_gaussian(k, sigma, width, fft_length);
    std::memcpy(kernel, k.data(), sizeof(float) * k.size());
    return BLUR_B200_OK;
}

int blur_b200_gauss_geometry(int rows, int cols, double sigma, int out[4])
{
    if (!out || rows < 1 || cols < 1) return BLUR_B200_EINVAL;
    const GaussGeometry g = gauss_geometry(rows, cols, sigma);
    out[0] = g.ksize; out[1] = g.pad; out[2] = g.n_col; out[3] = g.n_row;
    return BLUR_B200_OK;
}

int blur_b200_gauss_shard_geometry(int rows, int cols, double sigma, int nyquist_mode, int rank, int nranks,
                                   int64_t out[8])
{
    if (!out || rows < 1 || cols < 1 || !(sigma > 0) || nranks < 1 || rank < 0 || rank >= nranks) return BLUR_B200_EINVAL;
    const ShardGeom g = shard_geometry(gauss_geometry(rows, cols, sigma), nyquist_mode, rank, nranks);
    out[0] = g.algo; out[1] = g.stripe_rows; out[2] = g.cpp; out[3] = g.row0; out[4] = g.valid_rows;
    out[5] = std::max(0, std::min(2 * g.cpp, cols - rank * 2 * g.cpp));
    out[6] = static_cast<int64_t>(g.block_floats); out[7] = g.qpitch;
    return BLUR_B200_OK;
}

int blur_b200_gauss_plan_create(blur_b200_plan** plan, int rows, int cols, double sigma, int nyquist_mode, int device)
{
    if (!(sigma > 0)) return BLUR_B200_EINVAL;      // the negative encoding of the FFT box blur is internal (blur_b200_boxfft_*)
    return plan_create(plan, rows, cols, sigma, nyquist_mode, device, 0, 1);
}

int blur_b200_gauss_plan_create_sharded(blur_b200_plan** plan, int rows, int cols, double sigma, int nyquist_mode,
                                        int device, int rank, int nranks)
{
    if (!(sigma > 0)) return BLUR_B200_EINVAL;
    return plan_create(plan, rows, cols, sigma, nyquist_mode, device, rank, nranks);
}

void blur_b200_gauss_plan_destroy(blur_b200_plan* p)
{
    if (!p) return;
    cudaSetDevice(p->device);
    cudaFree(p->row.d_tw); cudaFree(p->row.d_kt);
    cudaFree(p->col.d_tw); cudaFree(p->col.d_kt);
    cudaFree(p->d_ols_twa_row); cudaFree(p->d_ols_twa_col); cudaFree(p->d_ols_twb);
    cudaFree(p->d_ols_kt_row); cudaFree(p->d_ols_kt_col);
    cudaFree(p->d_ols2_tw); cudaFree(p->d_ols2_kt);
    cudaFree(p->d_q); cudaFree(p->d_u); cudaFree(p->d_cnt); cudaFree(p->d_dtaps); cudaFree(p->d_corr_row); cudaFree(p->d_corr_col); cudaFree(p->d_insum); cudaFree(p->d_taps);
    for (uint8_t* b : p->d_sin) cudaFree(b);
    for (uint8_t* b : p->d_sout) cudaFree(b);
    if (p->h2d_stream) cudaStreamDestroy(p->h2d_stream);
    if (p->d2h_stream) cudaStreamDestroy(p->d2h_stream);
    for (cudaEvent_t e : p->ev_h2d) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : p->ev_out) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : p->ev_rowdone) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : p->ev_d2hdone) if (e) cudaEventDestroy(e);
    if (p->stream) cudaStreamDestroy(p->stream);
    for (cudaEvent_t e : p->ev_chunk) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : p->ev_stage) if (e) cudaEventDestroy(e);
    for (auto& row : p->ev_scopy) for (cudaEvent_t e : row) if (e) cudaEventDestroy(e);
    for (cudaStream_t cs : p->ce_stream) if (cs) cudaStreamDestroy(cs);
    for (cudaEvent_t e : p->ev_ce) if (e) cudaEventDestroy(e);
    delete p;
}

int blur_b200_gauss_plan_info(const blur_b200_plan* p, int64_t info[24])
{
    if (!p || !info) return BLUR_B200_EINVAL;
    info[0] = p->gg.ksize; info[1] = p->gg.pad; info[2] = p->gg.n_col; info[3] = p->gg.n_row;
    info[4] = p->stripe_rows; info[5] = p->cpp; info[6] = p->row0; info[7] = p->valid_rows;
    info[8] = static_cast<int64_t>(p->block_floats);
    info[9] = 3;
    info[10] = p->row.shape.threads; info[11] = p->col.shape.threads;
    info[12] = p->row.shape.batch; info[13] = p->col.shape.batch;
    info[14] = static_cast<int64_t>(p->row.shape.smem_bytes); info[15] = static_cast<int64_t>(p->col.shape.smem_bytes);
    info[16] = p->algo;
    info[17] = p->ols_row.L; info[18] = p->ols_col.L;
    info[19] = p->ols_row.V; info[20] = p->ols_col.V;
    info[21] = p->ols_row.nblocks; info[22] = p->ols_col.nblocks;
    info[23] = p->qpitch;
    if (p->algo == 1) {
        info[10] = info[11] = 64; info[12] = info[13] = 1;
        info[14] = 8 * p->ols_row.L + 16; info[15] = 8 * p->ols_col.L + 16;
    }
    return BLUR_B200_OK;
}

// Whole blur through the fused direct-convolution kernel (plans with p->direct).  The kernel's tiles read their
// neighbours' input pixels, so an in-place call goes through the plan's U buffer (at least one image large).
static int run_direct(blur_b200_plan* p, const uint8_t* d_in, uint8_t* d_out, cudaStream_t st)
{
    DirectArgs a{};
    a.in = d_in; a.rows = p->rows; a.cols = p->cols; a.pad = p->gg.pad; a.taps = p->d_dtaps;
    if (p->quirk_ols) {
        // pffft_ semantics: the same per-line Nyquist corrections the FFT epilogues add
        quirk_rows_kernel<<<p->rows, 128, 0, st>>>(d_in, p->cols, p->gg.pad, p->quirk_c_row, p->d_corr_row, p->rows,
                                                   p->cols % 4 == 0 && reinterpret_cast<uintptr_t>(d_in) % 4 == 0);
        CU_TRY(cudaGetLastError());
        ++g_launches;
        const int rc = quirk_prepare_cols(p, d_in, st);
        if (rc) return rc;
        a.corr_row = p->d_corr_row; a.corr_row_pitch = p->rows;
        a.corr_col = p->d_corr_col; a.corr_col_pitch = p->qpitch;
    }
    const size_t nbytes = static_cast<size_t>(p->rows) * p->cols * 3;
    const bool alias = d_in == d_out || (d_out < d_in + nbytes && d_in < d_out + nbytes);
    a.out = alias ? p->d_u : d_out;
    CU_TRY(launch_direct_gauss(a, st));
    ++g_launches;
    if (alias) CU_TRY(cudaMemcpyAsync(d_out, p->d_u, nbytes, cudaMemcpyDeviceToDevice, st));
    return BLUR_B200_OK;
}

int blur_b200_gauss_plan_uses_direct(const blur_b200_plan* p) { return p && p->direct ? 1 : 0; }

int blur_b200_gauss_plan_set_direct(blur_b200_plan* p, int enable)
{
    if (!p) return BLUR_B200_EINVAL;
    if (enable && !p->d_dtaps) return BLUR_B200_ESIZE;     // this plan was built without the direct kernel's taps
    p->direct = enable != 0;
    return BLUR_B200_OK;
}

int blur_b200_gauss_run_device(blur_b200_plan* p, const uint8_t* d_in, uint8_t* d_out, void* stream)
{
    if (!p || !d_in || !d_out || p->nranks != 1) return BLUR_B200_EINVAL;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (p->direct) return run_direct(p, d_in, d_out, st);
    int rc = run_row_pass(p, d_in, p->d_q, st);
    if (rc) return rc;
    bool fused = false;
    rc = run_col_pass_to_u(p, p->d_q, st, 0, -1, d_out, p->cols, p->cols, &fused);
    if (rc || fused) return rc;
    return run_interleave(p, d_out, p->cols, p->cols, st);
}

// Streamed host path: n host images through one plan.  Three streams — H2D, compute,
// D2H — and double-buffered device images: the upload of image i+1 and the download of image
// i-1 run underneath the kernels of image i, and inside one image the row pass starts on the
// first chunk of rows as soon as it has arrived while the column pass hands finished overlap-save
// blocks (V rows, all columns) to the download stream.  PCIe is full duplex, so the steady state
// costs one image transfer per direction per image instead of upload + blur + download.
int blur_b200_gauss_run_host_batch(blur_b200_plan* p, const uint8_t* const* h_in, uint8_t* const* h_out, int n)
{
    if (!p || !h_in || !h_out || n < 0 || p->nranks != 1) return BLUR_B200_EINVAL;
    for (int i = 0; i < n; ++i)
        if (!h_in[i] || !h_out[i]) return BLUR_B200_EINVAL;
    if (n == 0) return BLUR_B200_OK;
    CU_TRY(cudaSetDevice(p->device));
    const size_t pitch = static_cast<size_t>(p->cols) * 3;
    const size_t nbytes = pitch * p->rows;
    const int nbuf = n > 1 ? 2 : 1;
    if (!p->h2d_stream) {
        CU_TRY(cudaStreamCreateWithFlags(&p->h2d_stream, cudaStreamNonBlocking));
        CU_TRY(cudaStreamCreateWithFlags(&p->d2h_stream, cudaStreamNonBlocking));
        for (cudaEvent_t& e : p->ev_h2d) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (cudaEvent_t& e : p->ev_out) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (cudaEvent_t& e : p->ev_rowdone) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (cudaEvent_t& e : p->ev_d2hdone) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    for (int k = 0; k < nbuf; ++k) {
        if (!p->d_sin[k]) CU_TRY(cudaMalloc(&p->d_sin[k], nbytes));
        if (!p->d_sout[k]) CU_TRY(cudaMalloc(&p->d_sout[k], nbytes));
    }
    cudaStream_t sc = p->stream;
    const bool ols = p->algo == 1;
    // row chunks: whole waves of the persistent row grid (see blur_b200_gauss_row_pass_push)
    const int npairs = cdiv(p->rows, 2);
    int nchunks = 1, wave_pairs = npairs, nwaves = 1;
    if (ols) {
        wave_pairs = row_wave_pairs(p);
        nwaves = cdiv(npairs, wave_pairs);
        nchunks = std::max(1, std::min<int>(blur_b200_plan::kHostChunks, nwaves / 2));
    }
    const int nblk = ols ? std::min<int>(p->ols_col.nblocks, blur_b200_plan::kHostChunks) : 1;
    for (int i = 0; i < n; ++i) {
        const int k = i & (nbuf - 1);
        if (i >= nbuf) CU_TRY(cudaStreamWaitEvent(p->h2d_stream, p->ev_rowdone[k], 0));     // row pass of image i-2 has read d_sin[k]
        if (p->direct) {
            // narrow kernel: one upload, the fused direct kernel, one download; images still overlap on the three streams
            CU_TRY(cudaMemcpyAsync(p->d_sin[k], h_in[i], nbytes, cudaMemcpyHostToDevice, p->h2d_stream));
            CU_TRY(cudaEventRecord(p->ev_h2d[0], p->h2d_stream));
            CU_TRY(cudaStreamWaitEvent(sc, p->ev_h2d[0], 0));
            if (i >= nbuf) CU_TRY(cudaStreamWaitEvent(sc, p->ev_d2hdone[k], 0));             // image i-2 has left d_sout[k]
            const int rc = run_direct(p, p->d_sin[k], p->d_sout[k], sc);
            if (rc) return rc;
            CU_TRY(cudaEventRecord(p->ev_rowdone[k], sc));
            CU_TRY(cudaStreamWaitEvent(p->d2h_stream, p->ev_rowdone[k], 0));
            CU_TRY(cudaMemcpyAsync(h_out[i], p->d_sout[k], nbytes, cudaMemcpyDeviceToHost, p->d2h_stream));
            CU_TRY(cudaEventRecord(p->ev_d2hdone[k], p->d2h_stream));
            continue;
        }
        for (int c = 0; c < nchunks; ++c) {
            const int p0 = std::min(npairs, static_cast<int>(static_cast<int64_t>(nwaves) * c / nchunks) * wave_pairs);
            const int p1 = c + 1 == nchunks ? npairs : std::min(npairs, static_cast<int>(static_cast<int64_t>(nwaves) * (c + 1) / nchunks) * wave_pairs);
            const int r0 = 2 * p0, r1 = std::min(2 * p1, p->rows);
            if (r1 <= r0) continue;
            CU_TRY(cudaMemcpyAsync(p->d_sin[k] + r0 * pitch, h_in[i] + r0 * pitch, (r1 - r0) * pitch, cudaMemcpyHostToDevice, p->h2d_stream));
            CU_TRY(cudaEventRecord(p->ev_h2d[c], p->h2d_stream));
            CU_TRY(cudaStreamWaitEvent(sc, p->ev_h2d[c], 0));
            if (ols) {
                OlsRowIO io{};
                io.in = p->d_sin[k] + r0 * pitch;
                io.q = p->d_q + static_cast<size_t>(r0) * p->qpitch;
                io.rows = r1 - r0; io.cols = p->cols;
                io.npairs = cdiv(io.rows, 2);
                io.srows = p->stripe_rows; io.qpitch = p->qpitch; io.cpart = p->cpart; io.nparts = 1;
                io.ax = p->ols_row;
                io.nunits = io.ax.nblocks * io.npairs;
                if (p->quirk_ols) {
                    quirk_rows_kernel<<<io.rows, 128, 0, sc>>>(io.in, p->cols, p->gg.pad, p->quirk_c_row, p->d_corr_row + r0, p->rows,
                                                               p->cols % 4 == 0 && reinterpret_cast<uintptr_t>(io.in) % 4 == 0);
                    CU_TRY(cudaGetLastError());
                    ++g_launches;
                    io.corr = p->d_corr_row + r0; io.corr_pitch = p->rows;
                }
                CU_TRY(launch_row_fft(p, io, p->num_sms, sc));
                ++g_launches;
            } else {
                const int rc = run_row_pass(p, p->d_sin[k], p->d_q, sc);
                if (rc) return rc;
            }
        }
        if (p->quirk_ols) {
            const int rc = quirk_prepare_cols(p, p->d_sin[k], sc);
            if (rc) return rc;
        }
        CU_TRY(cudaEventRecord(p->ev_rowdone[k], sc));
        if (i >= nbuf) CU_TRY(cudaStreamWaitEvent(sc, p->ev_d2hdone[k], 0));                 // image i-2 has left d_sout[k]
        int rows_out = 0;
        for (int b = 0; b < nblk; ++b) {
            const bool last = b + 1 == nblk;
            bool fused = false;
            int rc = run_col_pass_to_u(p, p->d_q, sc, ols ? b : 0, ols ? (last ? p->ols_col.nblocks : b + 1) : -1, p->d_sout[k], p->cols,
                                       p->cols, &fused);
            if (rc) return rc;
            const int rend = last ? p->rows : std::min(p->rows, (b + 1) * p->ols_col.V) & ~3;
            if (rend <= rows_out) continue;
            if (!fused) {                                   // fused: the column kernel has written these rows of d_sout itself
                rc = run_interleave(p, p->d_sout[k], p->cols, p->cols, sc, rows_out, rend);
                if (rc) return rc;
            }
            CU_TRY(cudaEventRecord(p->ev_out[b], sc));
            CU_TRY(cudaStreamWaitEvent(p->d2h_stream, p->ev_out[b], 0));
            CU_TRY(cudaMemcpyAsync(h_out[i] + rows_out * pitch, p->d_sout[k] + rows_out * pitch, (rend - rows_out) * pitch,
                                   cudaMemcpyDeviceToHost, p->d2h_stream));
            rows_out = rend;
        }
        CU_TRY(cudaEventRecord(p->ev_d2hdone[k], p->d2h_stream));
    }
    CU_TRY(cudaStreamSynchronize(p->d2h_stream));
    CU_TRY(cudaStreamSynchronize(sc));
    return BLUR_B200_OK;
}

int blur_b200_gauss_run_host(blur_b200_plan* p, const uint8_t* h_in, uint8_t* h_out)
{
    return blur_b200_gauss_run_host_batch(p, &h_in, &h_out, 1);
}

// The one-shot entry keeps its last plan (tables, Q, U, staging images) so that a caller looping over
// same-sized images — the reference's `Test` usage — pays for the plan once; a different geometry
// replaces it.  Calls are serialised by the mutex.
static std::mutex g_oneshot_mu;
static blur_b200_plan* g_oneshot_plan = nullptr;

void blur_b200_release_cached(void)
{
    std::lock_guard<std::mutex> lock(g_oneshot_mu);
    if (g_oneshot_plan) blur_b200_gauss_plan_destroy(g_oneshot_plan);
    g_oneshot_plan = nullptr;
}

int blur_b200_gauss_u8c3(uint8_t* h_bgr, int rows, int cols, double sigma, int nyquist_mode)
{
    blur_b200_plan*& cached = g_oneshot_plan;
    std::lock_guard<std::mutex> lock(g_oneshot_mu);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { g_cuda_err = "cudaGetDevice failed (no CUDA device?)"; return BLUR_B200_ECUDA; }
    if (cached && (cached->rows != rows || cached->cols != cols || cached->sigma != sigma || cached->mode != nyquist_mode ||
                   cached->device != dev)) {
        blur_b200_gauss_plan_destroy(cached);
        cached = nullptr;
    }
    if (!cached) {
        const int rc = blur_b200_gauss_plan_create(&cached, rows, cols, sigma, nyquist_mode, dev);
        if (rc) { cached = nullptr; return rc; }
    }
    const int rc = blur_b200_gauss_run_host(cached, h_bgr, h_bgr);
    if (rc) { blur_b200_gauss_plan_destroy(cached); cached = nullptr; }
    return rc;
}

// ---- FFT box blur: the reference built with `#define boxblur` (Source.cpp:12) ---------------------
int blur_b200_boxfft_klen(int rows, int cols, double nsmooth)
{
    if (rows < 1 || cols < 1 || !(nsmooth >= 0)) return 0;
    return boxfft_klen(rows, cols, nsmooth);
}

int blur_b200_boxfft_plan_create(blur_b200_plan** plan, int rows, int cols, double nsmooth, int nyquist_mode, int device)
{
    if (!plan || rows < 1 || cols < 1 || !(nsmooth > 0)) return BLUR_B200_EINVAL;
    const int klen = boxfft_klen(rows, cols, nsmooth);
    // even kLen: the reference's pad (kLen - 1) / 2 * 2 is one short of the tent's half width and its outermost
    // taps wrap around the tile — undefined there, refused here; kLen = 1 is the identity
    if (klen < 3 || klen % 2 == 0) return BLUR_B200_EINVAL;
    return plan_create(plan, rows, cols, -static_cast<double>(klen), nyquist_mode, device, 0, 1);
}

int blur_b200_boxfft_u8c3(uint8_t* h_bgr, int rows, int cols, double nsmooth, int nyquist_mode)
{
    if (!h_bgr || rows < 1 || cols < 1 || !(nsmooth > 0)) return BLUR_B200_EINVAL;
    const int klen = boxfft_klen(rows, cols, nsmooth);
    if (klen < 3 || klen % 2 == 0) return BLUR_B200_EINVAL;
    blur_b200_plan*& cached = g_oneshot_plan;
    std::lock_guard<std::mutex> lock(g_oneshot_mu);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { g_cuda_err = "cudaGetDevice failed (no CUDA device?)"; return BLUR_B200_ECUDA; }
    const double key = -static_cast<double>(klen);
    if (cached && (cached->rows != rows || cached->cols != cols || cached->sigma != key || cached->mode != nyquist_mode ||
                   cached->device != dev)) {
        blur_b200_gauss_plan_destroy(cached);
        cached = nullptr;
    }
    if (!cached) {
        const int rc = plan_create(&cached, rows, cols, key, nyquist_mode, dev, 0, 1);
        if (rc) { cached = nullptr; return rc; }
    }
    const int rc = blur_b200_gauss_run_host(cached, h_bgr, h_bgr);
    if (rc) { blur_b200_gauss_plan_destroy(cached); cached = nullptr; }
    return rc;
}

int blur_b200_gauss_row_pass(blur_b200_plan* p, const uint8_t* d_in_stripe, float* d_q, void* stream)
{
    if (!p || !d_in_stripe || !d_q) return BLUR_B200_EINVAL;
    CU_TRY(cudaSetDevice(p->device));
    return run_row_pass(p, d_in_stripe, d_q, static_cast<cudaStream_t>(stream));
}

int blur_b200_gauss_row_pass_peer(blur_b200_plan* p, const uint8_t* d_in_stripe, float* const* q_peers, int npeers, void* stream)
{
    if (!p || !d_in_stripe || !q_peers || npeers != p->nranks || npeers > OlsRowIO::kMaxPeers) return BLUR_B200_EINVAL;
    if (p->algo != 1) return BLUR_B200_ESIZE;      // only the overlap-save row kernel has the fused epilogue
    for (int k = 0; k < npeers; ++k)
        if (!q_peers[k]) return BLUR_B200_EINVAL;
    CU_TRY(cudaSetDevice(p->device));
    return run_row_pass(p, d_in_stripe, p->d_q, static_cast<cudaStream_t>(stream), q_peers);
}

int blur_b200_gauss_row_pass_push(blur_b200_plan* p, const uint8_t* d_in_stripe, float* const* q_peers, int npeers,
                                  int nchunks, void* stream)
{
    if (!p || !d_in_stripe || !q_peers || npeers != p->nranks || npeers > OlsRowIO::kMaxPeers) return BLUR_B200_EINVAL;
    if (p->algo != 1) return BLUR_B200_ESIZE;
    for (int k = 0; k < npeers; ++k)
        if (!q_peers[k]) return BLUR_B200_EINVAL;
    if (p->valid_rows <= 0) return BLUR_B200_OK;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int N = p->nranks, rank = p->rank;
    if (!p->ev_chunk[0]) {
        int prio_lo = 0, prio_hi = 0;
        CU_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        for (cudaEvent_t& e : p->ev_chunk) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (int d = 0; d + 1 < N; ++d) {
            if (!p->ce_stream[d]) CU_TRY(cudaStreamCreateWithPriority(&p->ce_stream[d], cudaStreamNonBlocking, prio_hi));
            CU_TRY(cudaEventCreateWithFlags(&p->ev_ce[d], cudaEventDisableTiming));
        }
    }
    const int npairs = cdiv(p->valid_rows, 2);
    nchunks = std::max(1, std::min({nchunks, static_cast<int>(blur_b200_plan::kMaxChunks), npairs}));
    const size_t blk = static_cast<size_t>(p->stripe_rows) * p->qpitch;        // floats per (channel, part) block
    // Row kernel in "peer" addressing (ols_core.cuh): part == rank lands in our own receive buffer as
    // block (ch, rank); every other part lands in the local send buffer d_q as block (ch, part).
    float* base[OlsRowIO::kMaxPeers];
    for (int part = 0; part < N; ++part)
        base[part] = part == rank ? q_peers[rank] : p->d_q + (static_cast<ptrdiff_t>(part) - rank) * static_cast<ptrdiff_t>(blk);
    // Chunks are whole waves of the persistent row grid (grid CTAs / blocks per line row pairs), so
    // cutting the stripe adds no partially filled wave except the stripe's own last one.
    const int wave_pairs = row_wave_pairs(p);
    const int nwaves = cdiv(npairs, wave_pairs);
    nchunks = std::min(nchunks, nwaves);
    for (int c = 0; c < nchunks; ++c) {
        const int p0 = std::min(npairs, static_cast<int>(static_cast<int64_t>(nwaves) * c / nchunks) * wave_pairs);
        const int p1 = c + 1 == nchunks ? npairs : std::min(npairs, static_cast<int>(static_cast<int64_t>(nwaves) * (c + 1) / nchunks) * wave_pairs);
        const int r0 = 2 * p0, r1 = std::min(2 * p1, p->valid_rows);
        if (r1 <= r0) continue;
        OlsRowIO io{};
        io.in = d_in_stripe + static_cast<size_t>(r0) * p->cols * 3;
        io.q = p->d_q;
        io.rows = r1 - r0; io.cols = p->cols;
        io.npairs = cdiv(io.rows, 2);
        io.srows = p->stripe_rows; io.qpitch = p->qpitch; io.cpart = p->cpart; io.nparts = N;
        io.ax = p->ols_row;
        io.nunits = io.ax.nblocks * io.npairs;
        io.peer = 1; io.rank = rank;
        if (p->quirk_ols) { io.corr = p->d_corr_row + p->row0 + r0; io.corr_pitch = p->rows; }
        for (int part = 0; part < N; ++part) io.qpeer[part] = base[part] + static_cast<size_t>(r0) * p->qpitch;
        static const int diag = std::getenv("BLUR_B200_PUSH_DIAG") ? std::atoi(std::getenv("BLUR_B200_PUSH_DIAG")) : 0;   // timing experiments only
        if (diag != 1) {
            static const int sms_free = std::getenv("BLUR_B200_ROW_SMS_FREE") ? std::atoi(std::getenv("BLUR_B200_ROW_SMS_FREE")) : 0;
            CU_TRY(launch_row_fft(p, io, p->num_sms - (N > 1 ? sms_free : 0), st));
            ++g_launches;
        }
        if (N == 1 || diag == 2) continue;
        CU_TRY(cudaEventRecord(p->ev_chunk[c], st));
        // the GPU's copy engines move the finished chunk: per destination the three channel blocks sit at a pitch of
        // N blocks on both sides, i.e. one 2-D peer copy (3 rows) per destination and chunk, each destination on its
        // own stream.  (An SM copy kernel with 128-bit peer stores reached the same ~700 GB/s alone but did not get
        // scheduled beside the resident persistent row grid — profiles/r01_exchange_modes.txt — and was removed.)
        const size_t width = static_cast<size_t>(r1 - r0) * p->qpitch * sizeof(float);
        const size_t pitch = static_cast<size_t>(N) * blk * sizeof(float);
        for (int d = 0; d + 1 < N; ++d) {                   // start with the next rank: spread the NVSwitch ports
            const int part = (rank + 1 + d) % N;
            cudaStream_t cs = p->ce_stream[d];
            CU_TRY(cudaStreamWaitEvent(cs, p->ev_chunk[c], 0));
            CU_TRY(cudaMemcpy2DAsync(q_peers[part] + static_cast<size_t>(rank) * blk + static_cast<size_t>(r0) * p->qpitch, pitch,
                                     p->d_q + static_cast<size_t>(part) * blk + static_cast<size_t>(r0) * p->qpitch, pitch, width, 3,
                                     cudaMemcpyDeviceToDevice, cs));
        }
    }
    if (N > 1) {
        for (int d = 0; d + 1 < N; ++d) {
            CU_TRY(cudaEventRecord(p->ev_ce[d], p->ce_stream[d]));
            CU_TRY(cudaStreamWaitEvent(st, p->ev_ce[d], 0));
        }
    }
    return BLUR_B200_OK;
}

// ---- row-stripe ("halo") sharding ---------------------------------------------------------------------
// The column pass is overlap-save too, so a rank that KEEPS its row stripe needs, besides its own row-pass
// result, only `halo` rows of Q from the stripe above and below: 2 * halo * cols * 12 bytes per GPU instead
// of rows * cols * 12 * (P - 1) / P^2 (65 MB instead of 352 MB at 16384^2, sigma = 50, P = 8), and the result
// stays row-sharded like the input.  Stripes are whole column-pass blocks (V rows each) of the single-GPU
// block grid, so every block is transformed with exactly the rows, tables and arithmetic of the single-GPU
// plan: bit-identical output.  These entries work on an ordinary single-GPU plan of the WHOLE image (exact
// mode, two-factor column kernel); rank / nranks only select the stripe.  Q_local: float [3][cap + 2 halo][qpitch],
// image row g of the stripe's band at local row g - row0 + halo.
namespace {

struct HaloGeom { int row0, valid_rows, cap, halo, b0, nbl, qrows; };

bool halo_geom(const blur_b200_plan* p, int rank, int nranks, HaloGeom& g)
{
    if (!p || p->nranks != 1 || p->algo != 1 || p->quirk_ols || !p->ols_col.R2) return false;     // banded Q/U: two-factor column kernel
    if (nranks < 1 || rank < 0 || rank >= nranks) return false;
    const int V = p->ols_col.V, nb = p->ols_col.nblocks;
    const int bpr = cdiv(nb, nranks);
    g.halo = p->ols_col.halo;
    g.cap = bpr * V;
    g.b0 = std::min(nb, rank * bpr);
    g.nbl = std::max(0, std::min(nb, (rank + 1) * bpr) - g.b0);
    g.row0 = std::min(p->rows, g.b0 * V);
    g.valid_rows = std::max(0, std::min(p->rows, (g.b0 + g.nbl) * V) - g.row0);
    g.qrows = g.cap + 2 * g.halo;
    return true;
}

}  // namespace

int blur_b200_gauss_halo_info(const blur_b200_plan* p, int rank, int nranks, int64_t out[8])
{
    HaloGeom g;
    if (!out) return BLUR_B200_EINVAL;
    if (!halo_geom(p, rank, nranks, g)) return p && p->nranks == 1 ? BLUR_B200_ESIZE : BLUR_B200_EINVAL;
    out[0] = g.row0; out[1] = g.valid_rows; out[2] = g.cap; out[3] = g.halo; out[4] = p->qpitch;
    out[5] = static_cast<int64_t>(3) * g.qrows * p->qpitch; out[6] = g.b0; out[7] = g.nbl;
    return BLUR_B200_OK;
}

int blur_b200_gauss_halo_row_pass(blur_b200_plan* p, int rank, int nranks, const uint8_t* d_in_stripe, float* d_q_local, void* stream)
{
    HaloGeom g;
    if (!halo_geom(p, rank, nranks, g) || !d_q_local) return BLUR_B200_EINVAL;
    if (g.valid_rows <= 0) return BLUR_B200_OK;
    if (!d_in_stripe) return BLUR_B200_EINVAL;
    CU_TRY(cudaSetDevice(p->device));
    OlsRowIO io{};
    io.in = d_in_stripe; io.q = d_q_local + static_cast<size_t>(g.halo) * p->qpitch;
    io.rows = g.valid_rows; io.cols = p->cols;
    io.npairs = cdiv(g.valid_rows, 2);
    io.srows = g.qrows; io.qpitch = p->qpitch; io.cpart = p->cpart; io.nparts = 1;     // channel planes are qrows rows apart
    io.ax = p->ols_row;
    io.nunits = io.ax.nblocks * io.npairs;
    CU_TRY(launch_row_fft(p, io, p->num_sms, static_cast<cudaStream_t>(stream)));
    ++g_launches;
    return BLUR_B200_OK;
}

// this rank's first / last `halo` rows into the neighbours' bands (q_up = rank - 1's Q_local, q_down = rank + 1's;
// nullptr at the image edges): two 2-D peer copies (three channel planes each) on the copy engines
int blur_b200_gauss_halo_push(blur_b200_plan* p, int rank, int nranks, const float* d_q_local, float* q_up, float* q_down, void* stream)
{
    HaloGeom g;
    if (!halo_geom(p, rank, nranks, g) || !d_q_local) return BLUR_B200_EINVAL;
    if (g.valid_rows <= 0) return BLUR_B200_OK;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t qp = p->qpitch, plane = static_cast<size_t>(g.qrows) * qp * sizeof(float);
    const int n = std::min(g.halo, g.valid_rows);
    const size_t width = static_cast<size_t>(n) * qp * sizeof(float);
    if (q_up && rank > 0)        // my rows [0, n) are rows [cap, cap + n) of the upper neighbour's stripe: its local rows halo + cap ...
        CU_TRY(cudaMemcpy2DAsync(q_up + static_cast<size_t>(g.halo + g.cap) * qp, plane, d_q_local + static_cast<size_t>(g.halo) * qp, plane,
                                 width, 3, cudaMemcpyDeviceToDevice, st));
    if (q_down && rank + 1 < nranks && g.valid_rows == g.cap)    // my last n rows sit right above the lower neighbour's row 0: its local rows [halo - n, halo)
        CU_TRY(cudaMemcpy2DAsync(q_down + static_cast<size_t>(g.halo - n) * qp, plane,
                                 d_q_local + static_cast<size_t>(g.halo + g.valid_rows - n) * qp, plane, width, 3, cudaMemcpyDeviceToDevice, st));
    return BLUR_B200_OK;
}

// column pass + interleave of this rank's blocks: d_out_rows uint8 [valid_rows][cols][3]
int blur_b200_gauss_halo_col_pass(blur_b200_plan* p, int rank, int nranks, const float* d_q_local, uint8_t* d_out_rows, void* stream)
{
    HaloGeom g;
    if (!halo_geom(p, rank, nranks, g) || !d_q_local || !d_out_rows) return BLUR_B200_EINVAL;
    if (g.valid_rows <= 0) return BLUR_B200_OK;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int noct = std::min(p->qpitch / 8, cdiv(p->cols, 8));
    OlsColIO io{};
    io.rows = p->rows;                                  // image rows: block geometry, reflect patch, last block's valid count
    io.noct = noct;
    io.ax = p->ols_col;
    io.b0 = g.b0; io.nbl = g.nbl;
    io.nunits = ols2_col_npairs(io.noct, io.nbl);
    // virtual bases: where image row 0 would be
    io.u = p->d_u - static_cast<ptrdiff_t>(g.row0) * 8;
    io.urows = g.cap;
    io.qplane = static_cast<long long>(g.qrows) * p->qpitch;
    io.qrow_lo = std::max(0, g.row0 - g.halo);
    io.qrow_hi = std::min(p->rows, g.row0 + g.valid_rows + g.halo);
    const float* q_virtual = d_q_local + (static_cast<ptrdiff_t>(g.halo) - g.row0) * static_cast<ptrdiff_t>(p->qpitch);
    CU_TRY(launch_ols2_col(io, q_virtual, p->qpitch, g.qrows, p->d_ols2_tw, p->d_ols2_kt, p->num_sms, st));
    ++g_launches;
    const int vec = (static_cast<size_t>(p->cols) * 3 % 16 == 0) && (reinterpret_cast<uintptr_t>(d_out_rows) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(p->d_u) % 32 == 0) && (g.cap % 2 == 0);     // 16-byte loads of 4 rows: even plane height
    dim3 grid(cdiv(cdiv(noct, 2), 32), cdiv(cdiv(g.valid_rows, 4), 8));
    interleave_oct_kernel<<<grid, dim3(32, 8), 0, st>>>(p->d_u, d_out_rows, g.cap, noct, std::min(p->cols, noct * 8), p->cols, vec, 0, g.valid_rows);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

// ---- pffft Nyquist quirk of a SHARDED overlap-save plan (Source.cpp:414-427) --------------------------
// Row corrections are local to a stripe; the column corrections need sums over ALL rows: the integer column sums of
// the input (d_insum partials) and the row corrections themselves (d_corr_row).  Sequence per blur on every rank:
//   blur_b200_gauss_quirk_local  ->  row pass (+ exchange)  ->  caller sums both buffers over the ranks
//   (all-reduce SUM; every element of d_corr_row is non-zero on one rank only, d_insum is integer: exact)
//   ->  blur_b200_gauss_quirk_finish  ->  column pass.
int blur_b200_gauss_quirk_buffers(blur_b200_plan* p, float** d_corr_row, int64_t* n_corr_row, int** d_insum, int64_t* n_insum)
{
    if (!p || !d_corr_row || !n_corr_row || !d_insum || !n_insum) return BLUR_B200_EINVAL;
    if (!p->quirk_ols) return BLUR_B200_ESIZE;
    *d_corr_row = p->d_corr_row; *n_corr_row = static_cast<int64_t>(3) * p->rows;
    *d_insum = p->d_insum; *n_insum = static_cast<int64_t>(quirk_nseg(p)) * 3 * p->cols;
    return BLUR_B200_OK;
}

int blur_b200_gauss_quirk_local(blur_b200_plan* p, const uint8_t* d_in_stripe, void* stream)
{
    if (!p || (!d_in_stripe && p->valid_rows > 0)) return BLUR_B200_EINVAL;
    if (!p->quirk_ols) return BLUR_B200_ESIZE;
    CU_TRY(cudaSetDevice(p->device));
    return quirk_local(p, d_in_stripe, static_cast<cudaStream_t>(stream));
}

int blur_b200_gauss_quirk_finish(blur_b200_plan* p, void* stream)
{
    if (!p) return BLUR_B200_EINVAL;
    if (!p->quirk_ols) return BLUR_B200_ESIZE;
    CU_TRY(cudaSetDevice(p->device));
    return quirk_finish(p, static_cast<cudaStream_t>(stream));
}

// ---- staged exchange ------------------------------------------------------------------------------
// The part of every rank (cpart columns) is cut into `nstages` sub-ranges of w = cpart / nstages columns.
// Stage k of the row pass transforms, for all rows of the stripe, the overlap-save blocks that touch
// sub-range k of any rank's part (and no earlier sub-range); as soon as its launch has finished the copy
// engines move sub-range k to its owners while the next stage's FFTs run, and the column pass of
// sub-range k starts when every rank's copies of stage k have landed — the exchange (the distributed
// flip_block, Source.cpp:540) is hidden behind both FFT passes instead of sitting between them.
// Buffers (send buffer of the plan and every rank's receive buffer): [stage][3][nranks][stripe_rows][w].
namespace {

int staged_check(const blur_b200_plan* p, int nstages)
{
    if (!p || nstages < 1 || nstages > blur_b200_plan::kMaxStages) return BLUR_B200_EINVAL;
    if (p->algo != 1 || !p->ols_row.R2 || !p->ols_col.R2) return BLUR_B200_ESIZE;
    if (p->cpart % (16 * nstages) != 0) return BLUR_B200_ESIZE;            // whole octet pairs per sub-range
    const int w = p->cpart / nstages;
    if (p->nranks * nstages > OlsRowIO::kMaxParts || p->ols_row.nblocks > OlsRowIO::kMaxSel) return BLUR_B200_ESIZE;
    // bulk epilogue of the row kernel: 16-byte aligned pieces, a block spans at most kOls2BulkPieces sub-ranges
    if (p->ols_row.V % 4 != 0 || p->cols % 4 != 0 || p->ols_row.V > 3 * w) return BLUR_B200_ESIZE;
    return BLUR_B200_OK;
}

// blocks of stage k: a block belongs to the earliest sub-range index (column / w) % nstages it touches
void stage_blocks(const blur_b200_plan* p, int nstages, int k, unsigned char* sel, int& n)
{
    const int w = p->cpart / nstages, V = p->ols_row.V;
    n = 0;
    for (int b = 0; b < p->ols_row.nblocks; ++b) {
        const int x0 = b * V, x1 = std::min(p->cols, x0 + V) - 1;
        const int v0 = x0 / w, v1 = x1 / w;
        const int st = (v1 - v0 + 1 >= nstages || v0 % nstages + (v1 - v0) >= nstages) ? 0 : v0 % nstages;
        if (st == k) sel[n++] = static_cast<unsigned char>(b);
    }
}

}  // namespace

int blur_b200_gauss_staged_supported(const blur_b200_plan* p, int nstages) { return staged_check(p, nstages); }

int blur_b200_gauss_row_pass_staged(blur_b200_plan* p, const uint8_t* d_in_stripe, float* const* q_peers, int npeers,
                                    int nstages, int store, int row_ctas_per_sm, void* stream)
{
    if (!p || !d_in_stripe || !q_peers || npeers != p->nranks || npeers > OlsRowIO::kMaxPeers) return BLUR_B200_EINVAL;
    int rc = staged_check(p, nstages);
    if (rc) return rc;
    for (int k = 0; k < npeers; ++k)
        if (!q_peers[k]) return BLUR_B200_EINVAL;
    if (p->valid_rows <= 0) return BLUR_B200_OK;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int N = p->nranks, rank = p->rank, S = nstages;
    if (!p->ev_stage[0]) {
        int prio_lo = 0, prio_hi = 0;
        CU_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        for (int k = 0; k < blur_b200_plan::kMaxStages; ++k) {
            CU_TRY(cudaEventCreateWithFlags(&p->ev_stage[k], cudaEventDisableTiming));
            for (int d = 0; d < 7; ++d) CU_TRY(cudaEventCreateWithFlags(&p->ev_scopy[k][d], cudaEventDisableTiming));
        }
        for (int d = 0; d + 1 < N; ++d)
            if (!p->ce_stream[d]) CU_TRY(cudaStreamCreateWithPriority(&p->ce_stream[d], cudaStreamNonBlocking, prio_hi));
    }
    const int w = p->cpart / S;
    const size_t blk = static_cast<size_t>(p->stripe_rows) * w;            // floats per (stage, channel, source) block
    const size_t plane = static_cast<size_t>(3) * N * blk;                 // floats per stage
    p->staged_store = store != 0;
    // store mode: the launch may be held to fewer CTAs per SM so that the column pass of the previous sub-range
    // (another stream) finds room beside it while this launch waits on NVLink
    int row_sms = p->num_sms;
    if (row_ctas_per_sm > 0) row_sms = std::max(1, p->num_sms * std::min(row_ctas_per_sm, ols2_row_ctas_per_sm()) / ols2_row_ctas_per_sm());
    for (int k = 0; k < S; ++k) {
        OlsRowIO io{};
        io.in = d_in_stripe; io.q = p->d_q;
        io.rows = p->valid_rows; io.cols = p->cols;
        io.npairs = cdiv(p->valid_rows, 2);
        io.srows = p->stripe_rows; io.qpitch = w; io.cpart = w; io.nparts = N * S; io.chparts = N;
        io.ax = p->ols_row;
        io.peer = 1; io.rank = rank;
        if (p->quirk_ols) { io.corr = p->d_corr_row + p->row0; io.corr_pitch = p->rows; }
        // virtual part v = (destination d, sub-range kk): own part straight into the own receive buffer, the others
        // into the send buffer at the source-block slot of the destination (the kernel adds rank * blk)
        // (store mode: every part straight into its owner's receive buffer through bulk copies over NVLink)
        for (int d = 0; d < N; ++d)
            for (int kk = 0; kk < S; ++kk)
                io.qpeer[d * S + kk] = (d == rank || store) ? q_peers[d] + kk * plane
                                                            : p->d_q + kk * plane + (static_cast<ptrdiff_t>(d) - rank) * static_cast<ptrdiff_t>(blk);
        stage_blocks(p, S, k, io.bsel, io.nbsel);
        io.nunits = io.nbsel * io.npairs;
        if (io.nbsel > 0) {
            CU_TRY(launch_row_fft(p, io, row_sms, st));
            ++g_launches;
        }
        if (N == 1) continue;
        CU_TRY(cudaEventRecord(p->ev_stage[k], st));
        if (store) continue;
        const size_t width = static_cast<size_t>(p->valid_rows) * w * sizeof(float);
        const size_t pitch = static_cast<size_t>(N) * blk * sizeof(float);
        for (int di = 0; di + 1 < N; ++di) {                    // start with the next rank: spread the NVSwitch ports
            const int d = (rank + 1 + di) % N;
            cudaStream_t cs = p->ce_stream[di];
            CU_TRY(cudaStreamWaitEvent(cs, p->ev_stage[k], 0));
            CU_TRY(cudaMemcpy2DAsync(q_peers[d] + k * plane + static_cast<size_t>(rank) * blk, pitch,
                                     p->d_q + k * plane + static_cast<size_t>(d) * blk, pitch, width, 3, cudaMemcpyDeviceToDevice, cs));
            CU_TRY(cudaEventRecord(p->ev_scopy[k][di], cs));
        }
    }
    return BLUR_B200_OK;
}

int blur_b200_gauss_stage_wait(blur_b200_plan* p, int stage, void* stream)
{
    if (!p || stage < 0 || stage >= blur_b200_plan::kMaxStages) return BLUR_B200_EINVAL;
    if (p->valid_rows <= 0 || p->nranks == 1 || !p->ev_stage[0]) return BLUR_B200_OK;
    CU_TRY(cudaSetDevice(p->device));
    if (p->staged_store) {
        CU_TRY(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), p->ev_stage[stage], 0));
        return BLUR_B200_OK;
    }
    for (int di = 0; di + 1 < p->nranks; ++di) CU_TRY(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), p->ev_scopy[stage][di], 0));
    return BLUR_B200_OK;
}

int blur_b200_gauss_col_pass_stage(blur_b200_plan* p, const float* d_q_recv, uint8_t* d_out_cols, int stage, int nstages,
                                   int col_ctas_per_sm, void* stream)
{
    if (!p || !d_q_recv || !d_out_cols || stage < 0 || stage >= nstages) return BLUR_B200_EINVAL;
    int rc = staged_check(p, nstages);
    if (rc) return rc;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int N = p->nranks, S = nstages, w = p->cpart / S;
    const int noct_total = std::min(p->qpitch / 8, cdiv(std::max(0, p->cols - p->rank * p->cpart), 8));
    const int noct = std::max(0, std::min(w / 8, noct_total - stage * (w / 8)));
    if (noct <= 0) return BLUR_B200_OK;
    const size_t blk = static_cast<size_t>(p->stripe_rows) * w;
    const size_t plane = static_cast<size_t>(3) * N * blk;
    uint8_t* u = p->d_u + static_cast<size_t>(stage) * 3 * (w / 8) * p->rows * 8;       // [3][noct][rows][8] of this sub-range
    OlsColIO io{};
    io.u = u;
    io.rows = p->rows;
    io.noct = noct;
    io.ax = p->ols_col;
    io.b0 = 0;
    io.nbl = io.ax.nblocks;
    io.nunits = ols2_col_npairs(io.noct, io.nbl);
    if (p->quirk_ols) { io.corr = p->d_corr_col + stage * w; io.corr_pitch = p->qpitch; }
    int col_sms = p->num_sms;
    if (col_ctas_per_sm > 0) col_sms = std::max(1, p->num_sms * std::min(col_ctas_per_sm, ols2_col_ctas_per_sm()) / ols2_col_ctas_per_sm());
    const int out_cols = 2 * p->cpp;
    const int valid_cols = std::max(0, std::min(p->cpart, p->cols - p->rank * p->cpart));
    const int ncols = std::max(0, std::min(w, valid_cols - stage * w));
    uint8_t* out = d_out_cols + static_cast<size_t>(stage) * w * 3;
    const bool fuse = col_fuse_enabled(p) && ncols > 0;
    if (fuse) { io.out = out; io.out_cols = out_cols; io.ncols = std::min(ncols, noct * 8); io.cnt = p->d_cnt; }
    CU_TRY(launch_ols2_col(io, d_q_recv + stage * plane, w, N * p->stripe_rows, p->d_ols2_tw, p->d_ols2_kt, col_sms, st));
    ++g_launches;
    if (ncols <= 0 || fuse) return BLUR_B200_OK;
    const int vec = (static_cast<size_t>(out_cols) * 3 % 16 == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(u) % 32 == 0) && (p->rows % 4 == 0);
    dim3 g(cdiv(cdiv(noct, 2), 32), cdiv(cdiv(p->rows, 4), 8));
    interleave_oct_kernel<<<g, dim3(32, 8), 0, st>>>(u, out, p->rows, noct, std::min(ncols, noct * 8), out_cols, vec, 0, p->rows);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

int blur_b200_gauss_col_pass(blur_b200_plan* p, const float* d_q_chunks, uint8_t* d_out_cols, void* stream)
{
    if (!p || !d_q_chunks || !d_out_cols) return BLUR_B200_EINVAL;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    bool fused = false;
    int rc = run_col_pass_to_u(p, d_q_chunks, st, 0, -1, d_out_cols, 2 * p->cpp, 2 * p->cpp, &fused);
    if (rc || fused) return rc;
    return run_interleave(p, d_out_cols, 2 * p->cpp, 2 * p->cpp, st);
}

int blur_b200_gauss_rowpass_f32(blur_b200_plan* p, const uint8_t* d_in, float* d_planar, void* stream)
{
    if (!p || !d_in || !d_planar || p->nranks != 1) return BLUR_B200_EINVAL;
    CU_TRY(cudaSetDevice(p->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = run_row_pass(p, d_in, p->d_q, st);
    if (rc) return rc;
    dim3 grid(cdiv(p->cols, 256), p->rows, 3);
    if (p->algo == 1)
        q_ols_to_planar_kernel<<<grid, 256, 0, st>>>(p->d_q, d_planar, p->rows, p->cols, p->qpitch, p->cpart,
                                                     p->stripe_rows, p->nranks);
    else
        q_to_planar_kernel<<<grid, 256, 0, st>>>(p->d_q, d_planar, p->rows, p->cols, p->cpp, p->stripe_rows);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

// ---- box blur

int blur_b200_fastboxblur_device(uint8_t* d_img, int w, int h, int channels, int ksize, int passes, int batch,
                                 uint8_t* d_scratch, void* stream)
{
    if (!d_img || !d_scratch || w < 1 || h < 1 || channels < 1 || channels > 4 || ksize < 1 || passes < 0 || batch < 1)
        return BLUR_B200_EINVAL;
    const int r = (ksize - 1) / 2;
    if (r > w - 1 || r > h - 1) return BLUR_B200_EPAD;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    DivK dk;
    dk.k = 2u * static_cast<uint32_t>(r) + 1u;
    dk.d = 2u * dk.k;
    dk.magic = static_cast<uint32_t>(((1ull << 32) + dk.d - 1) / dk.d);
    dk.use_magic = (511ull * dk.k) * dk.d < (1ull << 32) ? 1 : 0;
    const size_t img_bytes = static_cast<size_t>(w) * h * channels;
    const size_t smem_h = static_cast<size_t>(w + 2 * r) * channels * sizeof(uint32_t);
    if (smem_h > 200 * 1024) return BLUR_B200_ESIZE;
    if (smem_h > 48 * 1024)     // per-device function attribute: set on every call, never cached process-wide
        CU_TRY(cudaFuncSetAttribute(boxblur_h_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_h)));
    // Fused path: one kernel per 2-D pass over the whole batch, ping-pong between image and scratch.
    if (!std::getenv("BLUR_B200_BOX_UNFUSED")) {
        uint8_t *src = d_img, *dst = d_scratch;
        bool fused_ok = true;
        for (int p = 0; p < passes; ++p) {
            const cudaError_t e = blurb200::launch_box_fused_pass(src, dst, w, h, channels, r, batch, img_bytes, st);
            if (e == cudaErrorNotSupported && p == 0) { fused_ok = false; break; }
            CU_TRY(e);
            ++g_launches;
            std::swap(src, dst);
        }
        if (fused_ok) {
            if (src != d_img) CU_TRY(cudaMemcpyAsync(d_img, src, img_bytes * batch, cudaMemcpyDeviceToDevice, st));
            return BLUR_B200_OK;
        }
    }
    const int pitch = w * channels;
    const bool vec = (pitch % 4 == 0) && (reinterpret_cast<uintptr_t>(d_img) % 4 == 0) &&
                     (reinterpret_cast<uintptr_t>(d_scratch) % 4 == 0) && (img_bytes % 4 == 0);
    const int seg = 64;
    // Keep each group's image + scratch resident in L2 across the 2*passes sweeps.
    const size_t l2_budget = 48ull << 20;
    int group = static_cast<int>(std::max<size_t>(1, l2_budget / (2 * img_bytes)));
    if (const char* e = std::getenv("BLUR_B200_BOX_GROUP")) group = std::max(1, std::atoi(e));
    group = std::min(group, 65535);
    for (int b0 = 0; b0 < batch; b0 += group) {
        const int nb = std::min(group, batch - b0);
        uint8_t* img = d_img + static_cast<size_t>(b0) * img_bytes;
        uint8_t* scr = d_scratch + static_cast<size_t>(b0) * img_bytes;
        for (int p = 0; p < passes; ++p) {
            boxblur_h_kernel<<<dim3(h, nb), 256, smem_h, st>>>(img, scr, w, h, channels, r, dk, img_bytes);
            CU_TRY(cudaGetLastError());
            if (vec)
                boxblur_v_kernel<4><<<dim3(cdiv(pitch / 4, 128), cdiv(h, seg), nb), 128, 0, st>>>(scr, img, pitch, h, r, dk, seg, img_bytes);
            else
                boxblur_v_kernel<1><<<dim3(cdiv(pitch, 128), cdiv(h, seg), nb), 128, 0, st>>>(scr, img, pitch, h, r, dk, seg, img_bytes);
            CU_TRY(cudaGetLastError());
            g_launches += 2;
        }
    }
    return BLUR_B200_OK;
}

int blur_b200_fastboxblur(uint8_t* h_img, int w, int h, int channels, int ksize, int passes)
{
    if (!h_img || w < 1 || h < 1 || channels < 1) return BLUR_B200_EINVAL;
    const size_t nbytes = static_cast<size_t>(w) * h * channels;
    uint8_t *d_img = nullptr, *d_scr = nullptr;
    CU_TRY(cudaMalloc(&d_img, nbytes));
    if (cudaMalloc(&d_scr, nbytes) != cudaSuccess) { cudaFree(d_img); g_cuda_err = "cudaMalloc scratch"; return BLUR_B200_ECUDA; }
    int rc = BLUR_B200_OK;
    if (cudaMemcpy(d_img, h_img, nbytes, cudaMemcpyHostToDevice) != cudaSuccess) rc = BLUR_B200_ECUDA;
    if (!rc) rc = blur_b200_fastboxblur_device(d_img, w, h, channels, ksize, passes, 1, d_scr, nullptr);
    if (!rc && cudaMemcpy(h_img, d_img, nbytes, cudaMemcpyDeviceToHost) != cudaSuccess) rc = BLUR_B200_ECUDA;
    cudaFree(d_img); cudaFree(d_scr);
    return rc;
}

// `batch` host images (contiguous, img bytes apart) streamed in groups: upload of group g+1, the
// kernels of group g and the download of group g-1 overlap on three streams (two device slots).
int blur_b200_fastboxblur_batch_host(uint8_t* h_imgs, int w, int h, int channels, int ksize, int passes, int batch)
{
    if (!h_imgs || w < 1 || h < 1 || channels < 1 || batch < 1) return BLUR_B200_EINVAL;
    const size_t img_bytes = static_cast<size_t>(w) * h * channels;
    const int group = static_cast<int>(std::max<size_t>(1, std::min<size_t>(batch, (256ull << 20) / img_bytes)));
    uint8_t *d_img[2] = {}, *d_scr = nullptr;
    cudaStream_t sh = nullptr, sc = nullptr, sd = nullptr;
    cudaEvent_t up[2] = {}, done[2] = {}, down[2] = {};
    int rc = BLUR_B200_OK;
    auto cu = [&](cudaError_t e) { if (e != cudaSuccess && !rc) { g_cuda_err = cudaGetErrorString(e); rc = BLUR_B200_ECUDA; } return e == cudaSuccess; };
    cu(cudaMalloc(&d_img[0], img_bytes * group)); cu(cudaMalloc(&d_img[1], img_bytes * group)); cu(cudaMalloc(&d_scr, img_bytes * group));
    cu(cudaStreamCreateWithFlags(&sh, cudaStreamNonBlocking)); cu(cudaStreamCreateWithFlags(&sc, cudaStreamNonBlocking));
    cu(cudaStreamCreateWithFlags(&sd, cudaStreamNonBlocking));
    for (int k = 0; k < 2; ++k) {
        cu(cudaEventCreateWithFlags(&up[k], cudaEventDisableTiming)); cu(cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming));
        cu(cudaEventCreateWithFlags(&down[k], cudaEventDisableTiming));
    }
    for (int b0 = 0, g = 0; b0 < batch && !rc; b0 += group, ++g) {
        const int nb = std::min(group, batch - b0), k = g & 1;
        uint8_t* hp = h_imgs + static_cast<size_t>(b0) * img_bytes;
        if (g >= 2) cu(cudaStreamWaitEvent(sh, down[k], 0));            // slot k has been downloaded
        cu(cudaMemcpyAsync(d_img[k], hp, img_bytes * nb, cudaMemcpyHostToDevice, sh));
        cu(cudaEventRecord(up[k], sh));
        cu(cudaStreamWaitEvent(sc, up[k], 0));
        if (!rc) rc = blur_b200_fastboxblur_device(d_img[k], w, h, channels, ksize, passes, nb, d_scr, sc);
        cu(cudaEventRecord(done[k], sc));
        cu(cudaStreamWaitEvent(sd, done[k], 0));
        cu(cudaMemcpyAsync(hp, d_img[k], img_bytes * nb, cudaMemcpyDeviceToHost, sd));
        cu(cudaEventRecord(down[k], sd));
    }
    if (sd) cu(cudaStreamSynchronize(sd));
    if (sc) cu(cudaStreamSynchronize(sc));
    if (sh) cu(cudaStreamSynchronize(sh));
    for (int k = 0; k < 2; ++k) {
        cudaFree(d_img[k]);
        if (up[k]) cudaEventDestroy(up[k]);
        if (done[k]) cudaEventDestroy(done[k]);
        if (down[k]) cudaEventDestroy(down[k]);
    }
    cudaFree(d_scr);
    if (sh) cudaStreamDestroy(sh);
    if (sc) cudaStreamDestroy(sc);
    if (sd) cudaStreamDestroy(sd);
    return rc;
}

// ---- layout utilities

int blur_b200_reflect_101_u8(const uint8_t* d_in, uint8_t* d_out, int top, int bottom, int left, int right,
                             const int original_size[2], int channels, void* stream)
{
    if (!d_in || !d_out || !original_size || channels < 1 || top < 0 || bottom < 0 || left < 0 || right < 0)
        return BLUR_B200_EINVAL;
    const int rows = original_size[0], cols = original_size[1];
    if (rows < 1 || cols < 1) return BLUR_B200_EINVAL;
    top = std::min(top, rows - 1); bottom = std::min(bottom, rows - 1);      // Utils.hpp:217-220
    left = std::min(left, cols - 1); right = std::min(right, cols - 1);
    const int prow = rows + top + bottom, pcolb = (cols + left + right) * channels;
    reflect101_u8_kernel<<<dim3(cdiv(pcolb, 256), prow), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        d_in, d_out, rows, cols, channels, top, bottom, left, right);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

int blur_b200_deinterleave_u8_f32(const uint8_t* d_in, float* d_b, float* d_g, float* d_r, uint32_t total, void* stream)
{
    if (!d_in || !d_b || !d_g || !d_r) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    deinterleave_u8_f32_kernel<<<cdiv(static_cast<int>(total), 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_in, d_b, d_g, d_r, total);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

int blur_b200_interleave_f32_u8(const float* d_b, const float* d_g, const float* d_r, uint8_t* d_out, uint32_t total, void* stream)
{
    if (!d_out || !d_b || !d_g || !d_r) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    interleave_f32_u8_kernel<<<cdiv(static_cast<int>(total), 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_b, d_g, d_r, d_out, total);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

int blur_b200_flip_block_f32(const float* d_in, float* d_out, int w, int h, void* stream)
{
    if (!d_in || !d_out || w < 1 || h < 1) return BLUR_B200_EINVAL;
    transpose_f32_kernel<<<dim3(cdiv(w, 32), cdiv(h, 32)), dim3(32, 8), 0, static_cast<cudaStream_t>(stream)>>>(d_in, d_out, w, h);
    CU_TRY(cudaGetLastError());
    ++g_launches;
    return BLUR_B200_OK;
}

int blur_b200_reflect_101_u8_host(const uint8_t* h_in, uint8_t* h_out, int top, int bottom, int left, int right,
                                  const int original_size[2], int channels)
{
    if (!h_in || !h_out || !original_size || channels < 1 || top < 0 || bottom < 0 || left < 0 || right < 0)
        return BLUR_B200_EINVAL;
    const int rows = original_size[0], cols = original_size[1];
    if (rows < 1 || cols < 1) return BLUR_B200_EINVAL;
    const int t = std::min(top, rows - 1), b = std::min(bottom, rows - 1);
    const int l = std::min(left, cols - 1), r = std::min(right, cols - 1);
    const size_t nin = static_cast<size_t>(rows) * cols * channels;
    const size_t nout = static_cast<size_t>(rows + t + b) * (cols + l + r) * channels;
    uint8_t *d_in = nullptr, *d_out = nullptr;
    CU_TRY(cudaMalloc(&d_in, nin));
    if (cudaMalloc(&d_out, nout) != cudaSuccess) { cudaFree(d_in); g_cuda_err = "cudaMalloc"; return BLUR_B200_ECUDA; }
    int rc = BLUR_B200_OK;
    if (cudaMemcpy(d_in, h_in, nin, cudaMemcpyHostToDevice) != cudaSuccess) rc = BLUR_B200_ECUDA;
    if (!rc) rc = blur_b200_reflect_101_u8(d_in, d_out, top, bottom, left, right, original_size, channels, nullptr);
    if (!rc && cudaMemcpy(h_out, d_out, nout, cudaMemcpyDeviceToHost) != cudaSuccess) rc = BLUR_B200_ECUDA;
    cudaFree(d_in); cudaFree(d_out);
    return rc;
}

// Host-buffer forms of the layout helpers: the reference's own signatures take host pointers
// (Utils.hpp:159-210, flip_block call sites Source.cpp:367,384,540,562).
namespace {
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t n) { return cudaMalloc(&p, n ? n : 1) == cudaSuccess ? 0 : BLUR_B200_ECUDA; }
};
}  // namespace

int blur_b200_deinterleave_u8_f32_host(const uint8_t* h_in, float* h_b, float* h_g, float* h_r, uint32_t total)
{
    if (!h_in || !h_b || !h_g || !h_r) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    DevBuf in, pl;
    const size_t n = total;
    if (in.alloc(3 * n) || pl.alloc(3 * n * sizeof(float))) { g_cuda_err = "cudaMalloc"; return BLUR_B200_ECUDA; }
    float* f = static_cast<float*>(pl.p);
    CU_TRY(cudaMemcpy(in.p, h_in, 3 * n, cudaMemcpyHostToDevice));
    int rc = blur_b200_deinterleave_u8_f32(static_cast<const uint8_t*>(in.p), f, f + n, f + 2 * n, total, nullptr);
    if (rc) return rc;
    float* dst[3] = {h_b, h_g, h_r};
    for (int c = 0; c < 3; ++c) CU_TRY(cudaMemcpy(dst[c], f + c * n, n * sizeof(float), cudaMemcpyDeviceToHost));
    return BLUR_B200_OK;
}

int blur_b200_interleave_f32_u8_host(const float* h_b, const float* h_g, const float* h_r, uint8_t* h_out, uint32_t total)
{
    if (!h_out || !h_b || !h_g || !h_r) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    DevBuf out, pl;
    const size_t n = total;
    if (out.alloc(3 * n) || pl.alloc(3 * n * sizeof(float))) { g_cuda_err = "cudaMalloc"; return BLUR_B200_ECUDA; }
    float* f = static_cast<float*>(pl.p);
    const float* src[3] = {h_b, h_g, h_r};
    for (int c = 0; c < 3; ++c) CU_TRY(cudaMemcpy(f + c * n, src[c], n * sizeof(float), cudaMemcpyHostToDevice));
    int rc = blur_b200_interleave_f32_u8(f, f + n, f + 2 * n, static_cast<uint8_t*>(out.p), total, nullptr);
    if (rc) return rc;
    CU_TRY(cudaMemcpy(h_out, out.p, 3 * n, cudaMemcpyDeviceToHost));
    return BLUR_B200_OK;
}

int blur_b200_flip_block_f32_host(const float* h_in, float* h_out, int w, int h)
{
    if (!h_in || !h_out || w < 1 || h < 1) return BLUR_B200_EINVAL;
    DevBuf a, b;
    const size_t nbytes = static_cast<size_t>(w) * h * sizeof(float);
    if (a.alloc(nbytes) || b.alloc(nbytes)) { g_cuda_err = "cudaMalloc"; return BLUR_B200_ECUDA; }
    CU_TRY(cudaMemcpy(a.p, h_in, nbytes, cudaMemcpyHostToDevice));
    int rc = blur_b200_flip_block_f32(static_cast<const float*>(a.p), static_cast<float*>(b.p), w, h, nullptr);
    if (rc) return rc;
    CU_TRY(cudaMemcpy(h_out, b.p, nbytes, cudaMemcpyDeviceToHost));
    return BLUR_B200_OK;
}

uint64_t blur_b200_kernel_launches(void) { return g_launches.load(); }

}  // extern "C"
