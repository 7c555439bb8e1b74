// dft_image.cu — pocketfft_2D + `#define DFT_image` spectrum path (Source.cpp:143-277, block
// :240-252): whole-image reflect-101 padding, per-channel 2-D real FFT, fft-shifted
// 20*log10(|Re F| + 1e-5) image, crop back to the input size.
//
// Two kernels, both built from the same pass functions as the blur (fft_core.cuh):
//   spec_rows_kernel : two padded rows packed as one complex length-W FFT; the scrambled result is
//                      un-scrambled through shared memory, split into the two rows' half spectra
//                      and written as G[ch][row][k], k in [0, W/2]           (the r2c of :238)
//   spec_cols_kernel : complex length-H FFT down each of the W/2+1 columns of G, then the
//                      reference's shift / column-only mirror / log mapping straight from the
//                      registers that hold the radix-32 outputs              (:241-252, :268-276)
#include <cuda_runtime.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/blur_b200.h"
#include "fft_dispatch.cuh"
#include "fft_plan.hpp"

namespace blurb200 { __attribute__((visibility("hidden"))) void note_launches(uint64_t n); }   // blur_b200.cu: internal launch counter
using namespace blurb200;

namespace {

struct Geom2D {
    int rows, cols;
    int border[4];     // top bottom left right (Source.cpp:160-176)
    int H, W;          // sizes[0], sizes[1]
    int pad;
};

// Source.cpp:149-176
bool geom2d(int rows, int cols, double nsmooth, Geom2D& g)
{
    g.rows = rows; g.cols = cols;
    const int ksize = gaussian_window(nsmooth, std::max(rows, cols));
    g.pad = (ksize - 1) / 2;
    if (g.pad < 0) return false;
    for (int i = 0; i < 4; ++i) g.border[i] = g.pad;
    int sizes[2] = {rows + 2 * g.pad, cols + 2 * g.pad};
    for (int i = 0; i < 2; ++i) {
        if (!is_valid_size(sizes[i])) {
            const int ns = nearest_transform_size(sizes[i]);
            const int np = ns - sizes[i];
            sizes[i] = ns;
            g.border[i * 2 + 0] += np / 2;
            g.border[i * 2 + 1] += static_cast<int>(np / 2.f + 0.5f);
        }
    }
    g.H = sizes[0]; g.W = sizes[1];
    // Reflect_101 clamps each border to dim-1 (Utils.hpp:217-220) which would desynchronise the
    // reference's own geometry; refuse such inputs.
    return g.border[0] <= rows - 1 && g.border[1] <= rows - 1 && g.border[2] <= cols - 1 && g.border[3] <= cols - 1;
}

__device__ __forceinline__ int refl_src(int i, int lead, int L)
{
    int s = i - lead;
    if (s < 0) s = -s;
    if (s >= L) s = 2 * (L - 1) - s;
    return s;
}

// frequency of logical position 32u (add q * N/32 for position 32u + q)
__device__ __forceinline__ int freq_base(const FftGeom& g, int u)
{
    int k = 0, mul = 1;
    for (int i = 0; i + 1 < g.npass; ++i) {
        const int m32 = g.M[i] >> 5;
        const int q = u / m32;
        u -= q * m32;
        k += q * mul;
        mul *= g.radix[i];
    }
    return k;
}

struct RowsIO {
    const uint8_t* in;    // [rows][cols][3]
    int rows, cols, top, left, H, W;
    struct Ctx { const uint8_t *pa, *pb; };
    __device__ __forceinline__ Ctx begin(int unit) const
    {
        const int ch = unit / (H / 2), rp = unit - ch * (H / 2);
        Ctx c;
        c.pa = in + static_cast<size_t>(refl_src(2 * rp, top, rows)) * cols * 3 + ch;
        c.pb = in + static_cast<size_t>(refl_src(2 * rp + 1, top, rows)) * cols * 3 + ch;
        return c;
    }
    __device__ __forceinline__ void load(const Ctx& c, int n, float& re, float& im) const
    {
        const int x = refl_src(n, left, cols) * 3;
        re = static_cast<float>(c.pa[x]);
        im = static_cast<float>(c.pb[x]);
    }
    __device__ __forceinline__ void store(const Ctx&, int, float, float) const {}
};

// grid = 3 * H/2 CTAs, one packed row pair each; blockDim >= W/32.
__global__ void __launch_bounds__(kMaxThreads, 1)
spec_rows_kernel(const __grid_constant__ FftGeom g, const __grid_constant__ RowsIO io,
                 const float2* __restrict__ tw, float2* __restrict__ G)
{
    extern __shared__ float2 smem[];
    const int T = blockDim.x, tid = threadIdx.x;
    const int last = g.npass - 1;
    const int unit = blockIdx.x;
    {
        const int W = g.W[0];
#define BLUR_F(R) for (int u = tid; u < W; u += T) fwd_pass<R, true>(smem, g, 0, u, tw, io, unit);
        BLUR_RADIX_SWITCH(g.radix[0], BLUR_F)
#undef BLUR_F
    }
    __syncthreads();
    for (int p = 1; p < last; ++p) {
        const int W = g.W[p];
#define BLUR_F(R) for (int u = tid; u < W; u += T) fwd_pass<R, false>(smem, g, p, u, tw, io, unit);
        BLUR_RADIX_SWITCH(g.radix[p], BLUR_F)
#undef BLUR_F
        __syncthreads();
    }
    // radix-32 forward in registers, then scatter to natural frequency order (needs every
    // thread's inputs in registers before anyone overwrites shared memory)
    const int N = g.N, W32 = N / 32;
    float xr[32], xi[32];
    const bool active = tid < W32;
    if (active) {
#pragma unroll
        for (int j = 0; j < 32; ++j) { const float2 v = smem[tid * 33 + j]; xr[j] = v.x; xi[j] = v.y; }
        Dft<32, false>::run(xr, xi);
    }
    __syncthreads();
    if (active) {
        const int kb = freq_base(g, tid);
#pragma unroll
        for (int q = 0; q < 32; ++q) {
            const int k = kb + q * W32;
            smem[k + (k >> 5)] = make_float2(xr[q], xi[q]);
        }
    }
    __syncthreads();
    // split Z = A + iB into the two real rows' spectra: A[k] = (Z[k] + conj Z[N-k]) / 2,
    // B[k] = (Z[k] - conj Z[N-k]) / (2i)
    const int H2 = io.H / 2, WC = N / 2 + 1;
    const int ch = unit / H2, rp = unit - ch * H2;
    // G is tiled by column quads: element (ch, row, k) lives at ((ch*NG + k/4)*H + row)*4 + k%4, so
    // four consecutive k of a row are one 32-byte sector here and one sector per row in the column
    // kernel (a plain [row][k] layout costs the column pass a whole sector per 8-byte element)
    const int NG = (WC + 3) >> 2;
    float2* ga = G + (static_cast<size_t>(ch) * NG * io.H + 2 * rp) * 4;
    for (int k = tid; k < WC; k += T) {
        const int nk = k ? N - k : 0;
        const float2 z = smem[k + (k >> 5)], c = smem[nk + (nk >> 5)];
        float2* d = ga + static_cast<size_t>(k >> 2) * io.H * 4 + (k & 3);
        d[0] = make_float2(0.5f * (z.x + c.x), 0.5f * (z.y - c.y));
        d[4] = make_float2(0.5f * (z.y + c.y), -0.5f * (z.x - c.x));
    }
}

struct ColsIO {
    const float2* G;   // [3][NG][H][4], NG = ceil(WC / 4) column quads
    int H, WC;
    struct Ctx { const float2* p; };
    __device__ __forceinline__ Ctx begin(int unit) const
    {
        const int ch = unit / WC, k = unit - ch * WC;
        Ctx c;
        c.p = G + (static_cast<size_t>(ch) * ((WC + 3) >> 2) + (k >> 2)) * H * 4 + (k & 3);
        return c;
    }
    __device__ __forceinline__ void load(const Ctx& c, int n, float& re, float& im) const
    {
        const float2 v = c.p[static_cast<size_t>(n) * 4];
        re = v.x; im = v.y;
    }
    __device__ __forceinline__ void store(const Ctx&, int, float, float) const {}
};

__device__ __forceinline__ uint8_t spec_to_u8(float v)
{
    const float t = v + 0.5f;                     // interleave_BGR, Utils.hpp:204-206
    return t > 0.f ? static_cast<uint8_t>(static_cast<int>(t)) : static_cast<uint8_t>(0);
}

// grid = 3 * (W/2+1) CTAs, one spectrum column each; blockDim >= H/32.
__global__ void __launch_bounds__(kMaxThreads, 1)
spec_cols_kernel(const __grid_constant__ FftGeom g, const __grid_constant__ ColsIO io,
                 const float2* __restrict__ tw, float* __restrict__ spec /* [3][H][W] or null */,
                 uint8_t* __restrict__ out /* [rows][cols][3] */, int W, int rows, int cols, int top, int left,
                 uint8_t* __restrict__ Tq /* [3][WC][H] staging of the fast path, or null */)
{
    extern __shared__ float2 smem[];
    const int T = blockDim.x, tid = threadIdx.x;
    const int last = g.npass - 1;
    const int unit = blockIdx.x;
    {
        const int Wk = g.W[0];
#define BLUR_F(R) for (int u = tid; u < Wk; u += T) fwd_pass<R, true>(smem, g, 0, u, tw, io, unit);
        BLUR_RADIX_SWITCH(g.radix[0], BLUR_F)
#undef BLUR_F
    }
    __syncthreads();
    for (int p = 1; p < last; ++p) {
        const int Wk = g.W[p];
#define BLUR_F(R) for (int u = tid; u < Wk; u += T) fwd_pass<R, false>(smem, g, p, u, tw, io, unit);
        BLUR_RADIX_SWITCH(g.radix[p], BLUR_F)
#undef BLUR_F
        __syncthreads();
    }
    const int H = g.N, H32 = H / 32;
    const int ch = unit / io.WC, k = unit - ch * io.WC;
    // output columns that read this spectrum column (Source.cpp:245-247): col_ = k, and for
    // k in [1, W/2-1] also col_ = W - k (cval = W/2 - col_ % (W/2)); col = col_ shifted by W/2
    const int colA = (k + W / 2) % W;
    const int colB = (k >= 1 && k <= W / 2 - 1) ? (W - k + W / 2) % W : -1;
    if (Tq && !spec) {
        // Fast path (uint8 image only): the column's H bytes are staged in shared memory in frequency
        // order and leave as one contiguous run of Tq[ch][k][m]; spec_finish_kernel applies the
        // shift / mirror / crop.  (Writing out[] from here costs one scattered byte store per pixel.)
        float xr[32], xi[32];
        const bool active = tid < H32;                     // blockDim >= H/32
        if (active) {
#pragma unroll
            for (int j = 0; j < 32; ++j) { const float2 v = smem[tid * 33 + j]; xr[j] = v.x; xi[j] = v.y; }
        }
        __syncthreads();
        uint8_t* sb = reinterpret_cast<uint8_t*>(smem);
        if (active) {
            Dft<32, false>::run(xr, xi);
            const int mb = freq_base(g, tid);
#pragma unroll
            for (int q = 0; q < 32; ++q) sb[mb + q * H32] = spec_to_u8(20.f * log10f(fabsf(xr[q]) + 0.00001f));
        }
        __syncthreads();
        uint4* dst = reinterpret_cast<uint4*>(Tq + (static_cast<size_t>(ch) * io.WC + k) * H);
        const uint4* src = reinterpret_cast<const uint4*>(sb);
        for (int i = tid; i < H / 16; i += T) dst[i] = src[i];
        return;
    }
    for (int u = tid; u < H32; u += T) {
        float xr[32], xi[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) { const float2 v = smem[u * 33 + j]; xr[j] = v.x; xi[j] = v.y; }
        Dft<32, false>::run(xr, xi);
        const int mb = freq_base(g, u);
#pragma unroll
        for (int q = 0; q < 32; ++q) {
            const int m = mb + q * H32;                    // row_ (frequency row)
            const int row = (m + H / 2) % H;               // inverse of row_ = (row + H/2) % H
            const float v = 20.f * log10f(fabsf(xr[q]) + 0.00001f);
            if (spec) {
                float* s = spec + (static_cast<size_t>(ch) * H + row) * W;
                s[colA] = v;
                if (colB >= 0) s[colB] = v;
            }
            const int r = row - top;
            if (r >= 0 && r < rows) {
                const uint8_t b = spec_to_u8(v);
                uint8_t* o = out + static_cast<size_t>(r) * cols * 3 + ch;
                const int ca = colA - left, cb = colB - left;
                if (ca >= 0 && ca < cols) o[ca * 3] = b;
                if (colB >= 0 && cb >= 0 && cb < cols) o[cb * 3] = b;
            }
        }
    }
}

// Tq [3][WC][H] (uint8, frequency order) -> out [rows][cols][3]: the reference's shift, column-only
// mirror and crop (Source.cpp:244-247, 268-276).  block (32, 8), tile = 32 output columns x 32 rows,
// transposed through shared memory: reads run along m (contiguous in Tq), writes along the
// interleaved output row.
__global__ void __launch_bounds__(256)
spec_finish_kernel(const uint8_t* __restrict__ Tq, uint8_t* __restrict__ out, int H, int W, int WC, int rows, int cols,
                   int top, int left)
{
    __shared__ uint8_t tile[3][32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
    const int r = r0 + tx;
    const int m = (r + top + H / 2) % H;
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i;
        if (c < cols && r < rows) {
            const int col_ = (c + left + W / 2) % W;
            const int k = col_ <= W / 2 ? col_ : W - col_;
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) tile[ch][i][tx] = Tq[(static_cast<size_t>(ch) * WC + k) * H + m];
        }
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int rr = r0 + i, c = c0 + tx;
        if (rr < rows && c < cols) {
            uint8_t* o = out + (static_cast<size_t>(rr) * cols + c) * 3;
            o[0] = tile[0][tx][i]; o[1] = tile[1][tx][i]; o[2] = tile[2][tx][i];
        }
    }
}

thread_local std::string g_err;
#define CU_TRY(expr)                                                             \
    do {                                                                         \
        cudaError_t _e = (expr);                                                 \
        if (_e != cudaSuccess) {                                                 \
            g_err = std::string(#expr) + ": " + cudaGetErrorString(_e);          \
            std::fprintf(stderr, "blur_b200 dft_image: %s\n", g_err.c_str());    \
            return BLUR_B200_ECUDA;                                              \
        }                                                                        \
    } while (0)

// Twiddles and the half-spectrum workspace G are kept between calls (one geometry per device at a
// time): the reference re-plans on every call (pocketfft's plan cache aside), here a repeated
// call costs two kernel launches.
struct SpecWs {
    int device = -1, H = 0, W = 0;
    float2 *tww = nullptr, *twh = nullptr, *G = nullptr;
    uint8_t* Tq = nullptr;
};
std::mutex g_ws_mutex;
SpecWs g_ws;

int threads_for(int n32) { return std::max(128, std::min(kMaxThreads, ((n32 + 31) / 32) * 32)); }

}  // namespace

extern "C" {


int blur_b200_dft_image_sizes(int rows, int cols, double nsmooth, int padded[2], int border[4])
{
    if (rows < 1 || cols < 1 || !padded) return BLUR_B200_EINVAL;
    Geom2D g;
    if (!geom2d(rows, cols, nsmooth, g)) return BLUR_B200_EPAD;
    padded[0] = g.H; padded[1] = g.W;
    if (border) for (int i = 0; i < 4; ++i) border[i] = g.border[i];
    return BLUR_B200_OK;
}

int blur_b200_dft_image_device(const uint8_t* d_in, uint8_t* d_out, int rows, int cols, double nsmooth,
                               float* d_spec_f32, int padded[2], void* stream)
{
    if (!d_in || !d_out || rows < 1 || cols < 1) return BLUR_B200_EINVAL;
    Geom2D gm;
    if (!geom2d(rows, cols, nsmooth, gm)) return BLUR_B200_EPAD;
    if (padded) { padded[0] = gm.H; padded[1] = gm.W; }
    FftGeom gw, gh;
    if (!make_geom(gm.W, gw) || !make_geom(gm.H, gh)) return BLUR_B200_ESIZE;
    // the un-scramble step keeps one radix-32 group per thread in registers
    if (gm.W / 32 > kMaxThreads || gm.H / 32 > kMaxThreads) return BLUR_B200_ESIZE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    std::lock_guard<std::mutex> lock(g_ws_mutex);
    int device = 0;
    CU_TRY(cudaGetDevice(&device));
    const int WC = gm.W / 2 + 1;
    if (g_ws.device != device || g_ws.H != gm.H || g_ws.W != gm.W) {
        if (g_ws.device >= 0) {
            CU_TRY(cudaDeviceSynchronize());
            cudaFree(g_ws.tww); cudaFree(g_ws.twh); cudaFree(g_ws.G); cudaFree(g_ws.Tq);
            g_ws = SpecWs();
        }
        std::vector<float2> tw;
        make_twiddles(gm.W, tw);
        CU_TRY(cudaMalloc(&g_ws.tww, sizeof(float2) * tw.size()));
        CU_TRY(cudaMemcpy(g_ws.tww, tw.data(), sizeof(float2) * tw.size(), cudaMemcpyHostToDevice));
        make_twiddles(gm.H, tw);
        CU_TRY(cudaMalloc(&g_ws.twh, sizeof(float2) * tw.size()));
        CU_TRY(cudaMemcpy(g_ws.twh, tw.data(), sizeof(float2) * tw.size(), cudaMemcpyHostToDevice));
        CU_TRY(cudaMalloc(&g_ws.G, sizeof(float2) * static_cast<size_t>(3) * ((WC + 3) / 4) * 4 * gm.H));
        CU_TRY(cudaMalloc(&g_ws.Tq, static_cast<size_t>(3) * WC * gm.H));
        g_ws.device = device; g_ws.H = gm.H; g_ws.W = gm.W;
    }
    float2 *d_tww = g_ws.tww, *d_twh = g_ws.twh, *d_G = g_ws.G;

    const size_t smem_w = static_cast<size_t>(gw.slots) * 8, smem_h = static_cast<size_t>(gh.slots) * 8;
    CU_TRY(cudaFuncSetAttribute(spec_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_w)));
    CU_TRY(cudaFuncSetAttribute(spec_cols_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_h)));
    RowsIO rio{d_in, rows, cols, gm.border[0], gm.border[2], gm.H, gm.W};
    spec_rows_kernel<<<3 * gm.H / 2, threads_for(gm.W / 32), smem_w, st>>>(gw, rio, d_tww, d_G);
    CU_TRY(cudaGetLastError());
    ColsIO cio{d_G, gm.H, WC};
    spec_cols_kernel<<<3 * WC, threads_for(gm.H / 32), smem_h, st>>>(gh, cio, d_twh, d_spec_f32, d_out, gm.W, rows, cols,
                                                                     gm.border[0], gm.border[2], g_ws.Tq);
    CU_TRY(cudaGetLastError());
    if (!d_spec_f32) {
        spec_finish_kernel<<<dim3((cols + 31) / 32, (rows + 31) / 32), dim3(32, 8), 0, st>>>(g_ws.Tq, d_out, gm.H, gm.W, WC, rows, cols,
                                                                                        gm.border[0], gm.border[2]);
        CU_TRY(cudaGetLastError());
        blurb200::note_launches(1);
    }
    blurb200::note_launches(2);
    CU_TRY(cudaStreamSynchronize(st));      // the shared workspace is free again when the call returns
    return BLUR_B200_OK;
}

int blur_b200_dft_image_host(const uint8_t* h_in, uint8_t* h_out, int rows, int cols, double nsmooth,
                             float* h_spec_f32, int padded[2])
{
    if (!h_in || !h_out || rows < 1 || cols < 1) return BLUR_B200_EINVAL;
    Geom2D gm;
    if (!geom2d(rows, cols, nsmooth, gm)) return BLUR_B200_EPAD;
    const size_t nb = static_cast<size_t>(rows) * cols * 3, ns = static_cast<size_t>(3) * gm.H * gm.W;
    uint8_t *d_in = nullptr, *d_out = nullptr;
    float* d_spec = nullptr;
    CU_TRY(cudaMalloc(&d_in, nb));
    CU_TRY(cudaMalloc(&d_out, nb));
    if (h_spec_f32) CU_TRY(cudaMalloc(&d_spec, ns * sizeof(float)));
    CU_TRY(cudaMemcpy(d_in, h_in, nb, cudaMemcpyHostToDevice));
    int rc = blur_b200_dft_image_device(d_in, d_out, rows, cols, nsmooth, d_spec, padded, nullptr);
    if (!rc) {
        CU_TRY(cudaMemcpy(h_out, d_out, nb, cudaMemcpyDeviceToHost));
        if (h_spec_f32) CU_TRY(cudaMemcpy(h_spec_f32, d_spec, ns * sizeof(float), cudaMemcpyDeviceToHost));
    }
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_spec);
    return rc;
}

}  // extern "C"
