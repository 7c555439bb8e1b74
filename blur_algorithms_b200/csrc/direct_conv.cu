// direct_conv.cu — narrow kernels: the separable blur as ONE fused direct-convolution kernel.
//
// For a kernel of a few dozen taps (sigma <= ~8: BASELINE config 1, 1080p sigma = 5, 31 taps) the FFT path is the
// wrong tool: a 1080p blur is a single partial wave of block FFTs per pass plus three launches, i.e. latency, not
// throughput.  The convolution the reference computes per line (reflect-101 extension, the float taps of getGaussian,
// Source.cpp:75-102; row pass :520-537, column pass :546-560, interleave_BGR's +0.5f truncation Utils.hpp:204-206) is
// evaluated here directly in fp32, both axes in one kernel, the horizontal result living only in shared memory:
//
//   stage   the uint8 tile (TY + 2 pad) x (TX + 2 pad) pixels, reflect-101 resolved per byte, into shared memory
//   h pass  work item = (row PAIR, channel, 8 consecutive pixels): the two rows ride the two halves of packed fp32
//           registers (FFMA2), a sliding window of 8 + 2 pad inputs in registers, taps broadcast -> H (float, smem)
//   v pass  work item = (pair of adjacent values of a row, 8 consecutive rows): H is read as float2 down the column,
//           same packed arithmetic, + 0.5f, truncation, bytes straight to the interleaved output
//
// pffft_'s Nyquist corrections (blur_b200.cu, quirk_* kernels) are added where the FFT kernels add them: after the
// row convolution (c * A_row * (-1)^x) and after the column convolution (c * A_col * (-1)^y).
// The float32 intermediate of this path never reaches HBM: 6 B/px of compulsory traffic instead of 30.
#include <cuda_runtime.h>
#include <cstdint>

#include "cplx2.cuh"
#include "direct_conv.hpp"

namespace blurb200 {

namespace {

constexpr int kTX = 64;                 // tile width in pixels
constexpr int kTY = 64;                 // tile height in rows
constexpr int kTV = 3 * kTX;            // values (pixel x channel) per tile row
constexpr int kSeg = 8;                 // outputs per work item along the convolution axis
constexpr int kThreads = 256;

template <int PAD> constexpr int in_pitch() { return ((kTX + 2 * PAD) * 3 + 3) & ~3; }
template <int PAD> constexpr int smem_bytes() { return (kTY + 2 * PAD) * (in_pitch<PAD>() + kTV * 4) + 4 * (PAD + 1); }

__device__ __forceinline__ int refl_clamp(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return min(max(i, 0), n - 1);      // only for tile positions whose outputs are never stored
}

// two uint8 -> packed floats, exactly, without the conversion pipe: 0x4B000000 | b is 2^23 + b
__device__ __forceinline__ cplx u8x2_to_f32x2(uint32_t a, uint32_t b)
{
    const cplx v = cpack(__uint_as_float(0x4B000000u | a), __uint_as_float(0x4B000000u | b));
    return csub(v, cpack(8388608.f, 8388608.f));
}

template <int PAD>
__global__ void __launch_bounds__(kThreads, 2)
direct_gauss_kernel(const __grid_constant__ DirectArgs a)
{
    constexpr int NR = kTY + 2 * PAD;            // staged rows
    constexpr int NW = kSeg + 2 * PAD;           // window of one work item
    extern __shared__ __align__(16) uint8_t smem[];
    float* H = reinterpret_cast<float*>(smem);                                  // [NR][kTV]
    float* taps = H + NR * kTV;                                                 // [PAD + 1]
    uint8_t* tin = reinterpret_cast<uint8_t*>(taps + PAD + 1);                  // [NR][in_pitch]
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kTY;
    const int rows = a.rows, cols = a.cols;

    for (int k = tid; k <= PAD; k += kThreads) taps[k] = a.taps[k];
    // ---- stage: reflect-101 per pixel (Source.cpp:525-529 / :549-551 build the same extension per line)
    constexpr int WP = kTX + 2 * PAD;
    for (int idx = tid; idx < NR * WP; idx += kThreads) {
        const int ry = idx / WP, px = idx - ry * WP;
        const int gy = refl_clamp(y0 - PAD + ry, rows), gx = refl_clamp(x0 - PAD + px, cols);
        const uint8_t* s = a.in + (static_cast<size_t>(gy) * cols + gx) * 3;
        uint8_t* d = tin + ry * in_pitch<PAD>() + 3 * px;
        d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
    }
    __syncthreads();

    // ---- h pass: item = (row pair, 8-pixel segment, channel)
    constexpr int NITEM_H = (NR / 2) * (kTX / kSeg) * 3;
    for (int item = tid; item < NITEM_H; item += kThreads) {
        const int ch = item % 3, rest = item / 3;
        const int seg = rest % (kTX / kSeg), rp = rest / (kTX / kSeg);
        const uint8_t* ra = tin + (2 * rp) * in_pitch<PAD>() + 3 * (seg * kSeg) + ch;
        const uint8_t* rb = ra + in_pitch<PAD>();
        cplx w[NW];
#pragma unroll
        for (int k = 0; k < NW; ++k) w[k] = u8x2_to_f32x2(ra[3 * k], rb[3 * k]);
        cplx acc[kSeg];
#pragma unroll
        for (int j = 0; j < kSeg; ++j) acc[j] = cscale(w[j + PAD], taps[0]);
#pragma unroll
        for (int k = 1; k <= PAD; ++k) {
            const float t = taps[k];
#pragma unroll
            for (int j = 0; j < kSeg; ++j) acc[j] = cfmar(cfmar(acc[j], t, w[j + PAD - k]), t, w[j + PAD + k]);
        }
        if (a.corr_row) {
            // pffft Nyquist term of the row pass: + (-1)^x * c A_row of the (reflected) image row
            const int ga = refl_clamp(y0 - PAD + 2 * rp, rows), gb = refl_clamp(y0 - PAD + 2 * rp + 1, rows);
            const float ca = a.corr_row[static_cast<size_t>(ch) * a.corr_row_pitch + ga];
            const float cb = a.corr_row[static_cast<size_t>(ch) * a.corr_row_pitch + gb];
#pragma unroll
            for (int j = 0; j < kSeg; ++j) {
                const int gx = refl_clamp(x0 + seg * kSeg + j, cols);
                const float sg = (gx & 1) ? -1.f : 1.f;
                acc[j] = cadd(acc[j], cpack(sg * ca, sg * cb));
            }
        }
        float* ha = H + (2 * rp) * kTV + 3 * (seg * kSeg) + ch;
#pragma unroll
        for (int j = 0; j < kSeg; ++j) {
            ha[3 * j] = cre(acc[j]);
            ha[kTV + 3 * j] = cim(acc[j]);
        }
    }
    __syncthreads();

    // ---- v pass: item = (value pair, 8-row segment)
    constexpr int NITEM_V = (kTV / 2) * (kTY / kSeg);
    for (int item = tid; item < NITEM_V; item += kThreads) {
        const int vp = item % (kTV / 2), seg = item / (kTV / 2);
        const cplx* hc = reinterpret_cast<const cplx*>(H + (seg * kSeg) * kTV + 2 * vp);
        cplx w[NW];
#pragma unroll
        for (int k = 0; k < NW; ++k) w[k] = hc[k * (kTV / 2)];
        cplx acc[kSeg];
#pragma unroll
        for (int j = 0; j < kSeg; ++j) acc[j] = cscale(w[j + PAD], taps[0]);
#pragma unroll
        for (int k = 1; k <= PAD; ++k) {
            const float t = taps[k];
#pragma unroll
            for (int j = 0; j < kSeg; ++j) acc[j] = cfmar(cfmar(acc[j], t, w[j + PAD - k]), t, w[j + PAD + k]);
        }
        // values 2 vp, 2 vp + 1 of the tile row: pixel / channel of each
        const int v0 = 2 * vp, v1 = v0 + 1;
        const int px0 = v0 / 3, c0 = v0 - 3 * px0, px1 = v1 / 3, c1 = v1 - 3 * px1;
        const int gx0 = x0 + px0, gx1 = x0 + px1;
        float cc0 = 0.f, cc1 = 0.f;
        if (a.corr_col) {
            if (gx0 < cols) cc0 = a.corr_col[static_cast<size_t>(c0) * a.corr_col_pitch + gx0];
            if (gx1 < cols) cc1 = a.corr_col[static_cast<size_t>(c1) * a.corr_col_pitch + gx1];
        }
#pragma unroll
        for (int j = 0; j < kSeg; ++j) {
            const int gy = y0 + seg * kSeg + j;
            if (gy >= rows) break;
            float r0 = cre(acc[j]), r1 = cim(acc[j]);
            if (a.corr_col) {
                const float sg = (gy & 1) ? -1.f : 1.f;
                r0 += sg * cc0; r1 += sg * cc1;
            }
            // (uint8_t)(v + 0.5f): truncation (Utils.hpp:189,204-206); negatives (pffft corrections only) saturate to 0
            uint8_t* o = a.out + static_cast<size_t>(gy) * cols * 3 + static_cast<size_t>(x0) * 3 + v0;
            if (gx0 < cols) o[0] = static_cast<uint8_t>(min(__float2uint_rz(fmaxf(r0 + 0.5f, 0.f)), 255u));
            if (gx1 < cols) o[1] = static_cast<uint8_t>(min(__float2uint_rz(fmaxf(r1 + 0.5f, 0.f)), 255u));
        }
    }
}

template <int PAD>
cudaError_t launch_pad(const DirectArgs& a, cudaStream_t st)
{
    constexpr int smem = smem_bytes<PAD>();
    cudaError_t e = cudaFuncSetAttribute(direct_gauss_kernel<PAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    const dim3 grid((a.cols + kTX - 1) / kTX, (a.rows + kTY - 1) / kTY);
    direct_gauss_kernel<PAD><<<grid, kThreads, smem, st>>>(a);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_direct_gauss(const DirectArgs& a, cudaStream_t st)
{
    switch (a.pad) {
#define BLUR_DIRECT_CASE(P) case P: return launch_pad<P>(a, st);
        BLUR_DIRECT_CASE(1) BLUR_DIRECT_CASE(2) BLUR_DIRECT_CASE(3) BLUR_DIRECT_CASE(4) BLUR_DIRECT_CASE(5) BLUR_DIRECT_CASE(6)
        BLUR_DIRECT_CASE(7) BLUR_DIRECT_CASE(8) BLUR_DIRECT_CASE(9) BLUR_DIRECT_CASE(10) BLUR_DIRECT_CASE(11) BLUR_DIRECT_CASE(12)
        BLUR_DIRECT_CASE(13) BLUR_DIRECT_CASE(14) BLUR_DIRECT_CASE(15) BLUR_DIRECT_CASE(16) BLUR_DIRECT_CASE(17) BLUR_DIRECT_CASE(18)
        BLUR_DIRECT_CASE(19) BLUR_DIRECT_CASE(20)
#undef BLUR_DIRECT_CASE
    default: return cudaErrorNotSupported;
    }
}

}  // namespace blurb200
