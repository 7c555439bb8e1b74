// layout_typed.cu — the reference's layout templates for element types beyond <uint8_t, float>:
//   deinterleave_BGR<T,U>  (Utils.hpp:159-184)   out[c][x] = (U)(in[3x+c] + round)
//   interleave_BGR<T,U>    (Utils.hpp:186-210)   out[3x+c] = (T)(in[c][x] + round)
//   flip_block<T,C>        (call sites Source.cpp:367,384,540,562; FastBoxBlur submodule)
// `round` follows Utils.hpp:163 / :189 exactly: 0.5f only when the destination is integral and the
// source is not; the sum is formed in float (the reference's `x + round` promotes to float) and the
// conversion to an integral destination truncates toward zero.  Out-of-range values are undefined
// behaviour in the reference (the blur never produces them); here they saturate, which is what the
// author's ARM build does and what the DFT_image fixtures pin for negatives.
//
// Element types are passed as BLUR_B200_T_* codes through one C entry per function; the C++ host
// layer (include/blur_algorithms.hpp) maps template arguments to the codes.
#include <cuda_runtime.h>
#include <cstdint>
#include <string>

#include "../../include/blur_b200.h"

namespace blurb200 {
__attribute__((visibility("hidden"))) void note_launches(uint64_t n);
__attribute__((visibility("hidden"))) void set_cuda_error(const char* what, cudaError_t e);
}  // namespace blurb200

namespace {

template <class T> struct is_int { static constexpr bool value = true; };
template <> struct is_int<float> { static constexpr bool value = false; };

// (D)(v + round): Utils.hpp:163 / :189 make `round` a float constant, so the sum is formed in float for
// every type pair (8/16-bit integers survive it exactly; int32 rounds to float precision, as there).
template <class D, class S>
__device__ __forceinline__ D convert_elem(S v)
{
    constexpr float round = is_int<D>::value && !is_int<S>::value ? 0.5f : 0.f;
    const float f = static_cast<float>(v) + round;
    if constexpr (!is_int<D>::value) return f;
    else if constexpr (sizeof(D) == 1) return static_cast<D>(__float2uint_rz(fminf(fmaxf(f, 0.f), 255.f)));
    else if constexpr (sizeof(D) == 2) return static_cast<D>(__float2uint_rz(fminf(fmaxf(f, 0.f), 65535.f)));
    else return static_cast<D>(__float2int_rz(f));         // cvt.rzi saturates
}

// Four pixels (12 interleaved elements) per thread: every thread reads / writes whole 4-element
// groups, so uint8 traffic is 32-bit and float traffic 128-bit wide on the aligned bulk.
template <class T, class U>
__global__ void __launch_bounds__(256)
deinterleave_kernel(const T* __restrict__ in, U* __restrict__ b, U* __restrict__ g, U* __restrict__ r, uint32_t total)
{
    const uint32_t x0 = 4u * (blockIdx.x * blockDim.x + threadIdx.x);
    if (x0 >= total) return;
    const uint32_t n = total - x0 < 4u ? total - x0 : 4u;
    const T* p = in + 3 * static_cast<size_t>(x0);
    T v[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) v[k] = static_cast<uint32_t>(k) < 3u * n ? p[k] : T(0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (static_cast<uint32_t>(k) < n) {
            b[x0 + k] = convert_elem<U, T>(v[3 * k + 0]);
            g[x0 + k] = convert_elem<U, T>(v[3 * k + 1]);
            r[x0 + k] = convert_elem<U, T>(v[3 * k + 2]);
        }
    }
}

template <class T, class U>
__global__ void __launch_bounds__(256)
interleave_kernel(const U* __restrict__ b, const U* __restrict__ g, const U* __restrict__ r, T* __restrict__ out, uint32_t total)
{
    const uint32_t x0 = 4u * (blockIdx.x * blockDim.x + threadIdx.x);
    if (x0 >= total) return;
    const uint32_t n = total - x0 < 4u ? total - x0 : 4u;
    T* p = out + 3 * static_cast<size_t>(x0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (static_cast<uint32_t>(k) < n) {
            p[3 * k + 0] = convert_elem<T, U>(b[x0 + k]);
            p[3 * k + 1] = convert_elem<T, U>(g[x0 + k]);
            p[3 * k + 2] = convert_elem<T, U>(r[x0 + k]);
        }
    }
}

// flip_block<T,C>: in is h x w pixels of PB = C * sizeof(T) bytes, out is w x h.  32 x 32 pixel tiles
// through shared memory; PB is a template constant for the pixel sizes the reference family uses.
template <int PB>
__global__ void __launch_bounds__(256)
flip_pixels_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int w, int h)
{
    struct Px { uint8_t b[PB]; };
    __shared__ Px tile[32][33];
    const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
    const Px* src = reinterpret_cast<const Px*>(in);
    Px* dst = reinterpret_cast<Px*>(out);
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int x = x0 + threadIdx.x, y = y0 + j;
        if (x < w && y < h) tile[j][threadIdx.x] = src[static_cast<size_t>(y) * w + x];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int y = y0 + threadIdx.x, x = x0 + j;
        if (x < w && y < h) dst[static_cast<size_t>(x) * h + y] = tile[threadIdx.x][j];
    }
}

inline int elem_size(int t)
{
    switch (t) {
        case BLUR_B200_T_U8: return 1;
        case BLUR_B200_T_U16: return 2;
        case BLUR_B200_T_I32: return 4;
        case BLUR_B200_T_F32: return 4;
        default: return 0;
    }
}

template <class F>
int dispatch2(int a, int b, F&& f)
{
    auto second = [&](auto ta) {
        switch (b) {
            case BLUR_B200_T_U8: return f(ta, uint8_t{});
            case BLUR_B200_T_U16: return f(ta, uint16_t{});
            case BLUR_B200_T_I32: return f(ta, int32_t{});
            case BLUR_B200_T_F32: return f(ta, float{});
            default: return static_cast<int>(BLUR_B200_EINVAL);
        }
    };
    switch (a) {
        case BLUR_B200_T_U8: return second(uint8_t{});
        case BLUR_B200_T_U16: return second(uint16_t{});
        case BLUR_B200_T_I32: return second(int32_t{});
        case BLUR_B200_T_F32: return second(float{});
        default: return static_cast<int>(BLUR_B200_EINVAL);
    }
}

int finish_launch(const char* what)
{
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        blurb200::set_cuda_error(what, e);
        return BLUR_B200_ECUDA;
    }
    blurb200::note_launches(1);
    return BLUR_B200_OK;
}

struct DevMem {
    void* p = nullptr;
    ~DevMem() { if (p) cudaFree(p); }
    bool alloc(size_t n) { return cudaMalloc(&p, n ? n : 1) == cudaSuccess; }
};

#define HOST_TRY(expr, what)                                   \
    do {                                                       \
        const cudaError_t _e = (expr);                         \
        if (_e != cudaSuccess) {                               \
            blurb200::set_cuda_error(what, _e);                \
            return BLUR_B200_ECUDA;                            \
        }                                                      \
    } while (0)

}  // namespace

extern "C" {

int blur_b200_deinterleave_typed(const void* d_in, void* d_b, void* d_g, void* d_r, uint32_t total, int in_type, int out_type,
                                 void* stream)
{
    if (!d_in || !d_b || !d_g || !d_r || !elem_size(in_type) || !elem_size(out_type)) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    const unsigned grid = (total + 1023u) / 1024u;
    return dispatch2(in_type, out_type, [&](auto tt, auto uu) {
        using T = decltype(tt);
        using U = decltype(uu);
        deinterleave_kernel<T, U><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const T*>(d_in), static_cast<U*>(d_b), static_cast<U*>(d_g), static_cast<U*>(d_r), total);
        return finish_launch("deinterleave_kernel");
    });
}

int blur_b200_interleave_typed(const void* d_b, const void* d_g, const void* d_r, void* d_out, uint32_t total, int plane_type,
                               int out_type, void* stream)
{
    if (!d_out || !d_b || !d_g || !d_r || !elem_size(plane_type) || !elem_size(out_type)) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    const unsigned grid = (total + 1023u) / 1024u;
    return dispatch2(out_type, plane_type, [&](auto tt, auto uu) {
        using T = decltype(tt);
        using U = decltype(uu);
        interleave_kernel<T, U><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const U*>(d_b), static_cast<const U*>(d_g), static_cast<const U*>(d_r), static_cast<T*>(d_out), total);
        return finish_launch("interleave_kernel");
    });
}

int blur_b200_flip_block_pixels(const void* d_in, void* d_out, int w, int h, int pixel_bytes, void* stream)
{
    if (!d_in || !d_out || w < 1 || h < 1) return BLUR_B200_EINVAL;
    const dim3 grid((w + 31) / 32, (h + 31) / 32), block(32, 8);
    const uint8_t* in = static_cast<const uint8_t*>(d_in);
    uint8_t* out = static_cast<uint8_t*>(d_out);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (pixel_bytes) {
        case 1: flip_pixels_kernel<1><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 2: flip_pixels_kernel<2><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 3: flip_pixels_kernel<3><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 4: flip_pixels_kernel<4><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 6: flip_pixels_kernel<6><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 8: flip_pixels_kernel<8><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 12: flip_pixels_kernel<12><<<grid, block, 0, st>>>(in, out, w, h); break;
        case 16: flip_pixels_kernel<16><<<grid, block, 0, st>>>(in, out, w, h); break;
        default: return BLUR_B200_EINVAL;
    }
    return finish_launch("flip_pixels_kernel");
}

int blur_b200_deinterleave_typed_host(const void* h_in, void* h_b, void* h_g, void* h_r, uint32_t total, int in_type, int out_type)
{
    const int si = elem_size(in_type), so = elem_size(out_type);
    if (!h_in || !h_b || !h_g || !h_r || !si || !so) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    const size_t n = total;
    DevMem in, pl;
    if (!in.alloc(3 * n * si) || !pl.alloc(3 * n * so)) { blurb200::set_cuda_error("cudaMalloc", cudaErrorMemoryAllocation); return BLUR_B200_ECUDA; }
    HOST_TRY(cudaMemcpy(in.p, h_in, 3 * n * si, cudaMemcpyHostToDevice), "cudaMemcpy H2D");
    uint8_t* f = static_cast<uint8_t*>(pl.p);
    const int rc = blur_b200_deinterleave_typed(in.p, f, f + n * so, f + 2 * n * so, total, in_type, out_type, nullptr);
    if (rc) return rc;
    void* dst[3] = {h_b, h_g, h_r};
    for (int c = 0; c < 3; ++c) HOST_TRY(cudaMemcpy(dst[c], f + c * n * so, n * so, cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
    return BLUR_B200_OK;
}

int blur_b200_interleave_typed_host(const void* h_b, const void* h_g, const void* h_r, void* h_out, uint32_t total, int plane_type,
                                    int out_type)
{
    const int sp = elem_size(plane_type), so = elem_size(out_type);
    if (!h_out || !h_b || !h_g || !h_r || !sp || !so) return BLUR_B200_EINVAL;
    if (!total) return BLUR_B200_OK;
    const size_t n = total;
    DevMem out, pl;
    if (!out.alloc(3 * n * so) || !pl.alloc(3 * n * sp)) { blurb200::set_cuda_error("cudaMalloc", cudaErrorMemoryAllocation); return BLUR_B200_ECUDA; }
    uint8_t* f = static_cast<uint8_t*>(pl.p);
    const void* src[3] = {h_b, h_g, h_r};
    for (int c = 0; c < 3; ++c) HOST_TRY(cudaMemcpy(f + c * n * sp, src[c], n * sp, cudaMemcpyHostToDevice), "cudaMemcpy H2D");
    const int rc = blur_b200_interleave_typed(f, f + n * sp, f + 2 * n * sp, out.p, total, plane_type, out_type, nullptr);
    if (rc) return rc;
    HOST_TRY(cudaMemcpy(h_out, out.p, 3 * n * so, cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
    return BLUR_B200_OK;
}

int blur_b200_flip_block_pixels_host(const void* h_in, void* h_out, int w, int h, int pixel_bytes)
{
    if (!h_in || !h_out || w < 1 || h < 1 || pixel_bytes < 1) return BLUR_B200_EINVAL;
    const size_t nbytes = static_cast<size_t>(w) * h * pixel_bytes;
    DevMem a, b;
    if (!a.alloc(nbytes) || !b.alloc(nbytes)) { blurb200::set_cuda_error("cudaMalloc", cudaErrorMemoryAllocation); return BLUR_B200_ECUDA; }
    HOST_TRY(cudaMemcpy(a.p, h_in, nbytes, cudaMemcpyHostToDevice), "cudaMemcpy H2D");
    const int rc = blur_b200_flip_block_pixels(a.p, b.p, w, h, pixel_bytes, nullptr);
    if (rc) return rc;
    HOST_TRY(cudaMemcpy(h_out, b.p, nbytes, cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
    return BLUR_B200_OK;
}

}  // extern "C"
