// blur_algorithms.hpp — header-only C++ host layer over the C ABI (include/blur_b200.h).
//
// Same function names, argument order and meaning as the hot-path functions of
// michelerenzullo/Blur_algorithms (reference @ 2f873a4e), on raw buffers instead of cv::Mat
// (only image.data / size[0] / size[1] are touched by the reference, Source.cpp:429-461).  A
// maintainer swaps the bodies in Source.cpp for these calls: see INTEGRATION.md.
//
// No CUDA headers are needed to include this file; link with -lblur_b200.  Every function throws
// blur_b200::Error (status code + text) where the reference would have had undefined behaviour
// or where the GPU is missing: there is no CPU fallback.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "blur_b200.h"

namespace blur_b200 {

struct Error : std::runtime_error {
    int code;
    Error(int c, const char* where)
        : std::runtime_error(std::string(where) + ": " + blur_b200_strerror(c) +
                             (c == BLUR_B200_ECUDA ? std::string(" (") + blur_b200_last_cuda_error() + ")" : std::string())),
          code(c) {}
};
inline void check(int rc, const char* where) { if (rc) throw Error(rc, where); }

// RAII plan: the cached twiddles + separable Gaussian spectra + workspaces (replaces the per-call
// pffft_new_setup / getGaussian / kernel FFT of Source.cpp:465-503).
class GaussPlan {
public:
    GaussPlan(int rows, int cols, double sigma, int nyquist_mode = BLUR_B200_NYQUIST_EXACT, int device = 0)
    { check(blur_b200_gauss_plan_create(&p_, rows, cols, sigma, nyquist_mode, device), "gauss_plan_create"); }
    ~GaussPlan() { if (p_) blur_b200_gauss_plan_destroy(p_); }
    GaussPlan(const GaussPlan&) = delete;
    GaussPlan& operator=(const GaussPlan&) = delete;
    // host buffers, in place allowed (image.data)
    void run(const uint8_t* h_in, uint8_t* h_out) { check(blur_b200_gauss_run_host(p_, h_in, h_out), "gauss_run_host"); }
    // n host images streamed through the plan (uploads, kernels and downloads overlap)
    void run_batch(const uint8_t* const* h_in, uint8_t* const* h_out, int n)
    { check(blur_b200_gauss_run_host_batch(p_, h_in, h_out, n), "gauss_run_host_batch"); }
    // device buffers on the plan's device, in place allowed; stream = cudaStream_t
    void run_device(const uint8_t* d_in, uint8_t* d_out, void* stream = nullptr)
    { check(blur_b200_gauss_run_device(p_, d_in, d_out, stream), "gauss_run_device"); }
    blur_b200_plan* get() const { return p_; }
private:
    blur_b200_plan* p_ = nullptr;
};

}  // namespace blur_b200

// ---- the reference's names (global namespace, like Source.cpp / Utils.hpp) ----------------------

// Utils.hpp:141-157
inline int isValidSize(int N) { return blur_b200_is_valid_size(N); }
inline int nearestTransformSize(int N) { return blur_b200_nearest_transform_size(N); }

// Source.cpp:60-73
inline int gaussian_window(double sigma, int max_width = 0) { return blur_b200_gaussian_window(sigma, max_width); }

// Source.cpp:75-102: `kernel` is resized to FFT_length (or width) like the reference does
template <typename T>
inline void getGaussian(T& kernel, double sigma, int width = 0, int FFT_length = 0)
{
    if (!width) width = gaussian_window(sigma);
    const int n = FFT_length ? FFT_length : width;
    std::vector<float> tmp(static_cast<size_t>(n > width ? n : width));
    blur_b200::check(blur_b200_get_gaussian(tmp.data(), sigma, width, FFT_length), "getGaussian");
    kernel.resize(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) kernel[static_cast<size_t>(i)] = tmp[static_cast<size_t>(i)];
}

// Source.cpp:429 `void pffft_(cv::Mat& image, double nsmooth)`: image.data (uint8 BGR interleaved,
// rows x cols) blurred in place, pffft's Nyquist convention (Source.cpp:414-427).
inline void pffft_(uint8_t* image_data, int rows, int cols, double nsmooth)
{ blur_b200::check(blur_b200_gauss_u8c3(image_data, rows, cols, nsmooth, BLUR_B200_NYQUIST_PFFFT), "pffft_"); }

// Source.cpp:280 `void pocketfft_1D(cv::Mat& image, double nsmooth)`: exact multiply on every bin.
inline void pocketfft_1D(uint8_t* image_data, int rows, int cols, double nsmooth)
{ blur_b200::check(blur_b200_gauss_u8c3(image_data, rows, cols, nsmooth, BLUR_B200_NYQUIST_EXACT), "pocketfft_1D"); }

// Source.cpp:143 `void pocketfft_2D(cv::Mat& image, double nsmooth)` without DFT_image: the same
// function of the input as pocketfft_1D up to float rounding.
inline void pocketfft_2D(uint8_t* image_data, int rows, int cols, double nsmooth)
{ blur_b200::check(blur_b200_gauss_u8c3(image_data, rows, cols, nsmooth, BLUR_B200_NYQUIST_EXACT), "pocketfft_2D"); }

// pocketfft_2D compiled with `#define DFT_image` (Source.cpp:240-252,268-276): the image is
// replaced by its log-magnitude spectrum picture.
inline void pocketfft_2D_DFT_image(uint8_t* image_data, int rows, int cols, double nsmooth)
{
    int padded[2];
    blur_b200::check(blur_b200_dft_image_host(image_data, image_data, rows, cols, nsmooth, nullptr, padded), "DFT_image");
}

// The same three entry points of the reference built with `#define boxblur` (Source.cpp:12): FFT convolution with
// box_kernel's tent (Source.cpp:128-140) instead of the Gaussian.  pffft = true reproduces pffft_'s Nyquist quirk.
inline void boxblur_fft(uint8_t* image_data, int rows, int cols, double nsmooth, bool pffft = false)
{ blur_b200::check(blur_b200_boxfft_u8c3(image_data, rows, cols, nsmooth, pffft ? BLUR_B200_NYQUIST_PFFFT : BLUR_B200_NYQUIST_EXACT), "boxblur_fft"); }

// FastBoxBlur's entry as called at Source.cpp:587: in place.
inline void fastboxblur(uint8_t* in, int w, int h, int channels, int ksize, int passes)
{ blur_b200::check(blur_b200_fastboxblur(in, w, h, channels, ksize, passes), "fastboxblur"); }
// `batch` contiguous images, streamed (uploads / kernels / downloads overlap)
inline void fastboxblur_batch(uint8_t* in, int w, int h, int channels, int ksize, int passes, int batch)
{ blur_b200::check(blur_b200_fastboxblur_batch_host(in, w, h, channels, ksize, passes, batch), "fastboxblur_batch"); }

// Utils.hpp:212-243.  Any trivially copyable T: a pixel is C*sizeof(T) bytes and reflection only
// moves whole pixels.  original_size = {rows, cols}; pads are clamped to dim-1 like the reference.
template <typename T, int C>
inline void Reflect_101(const T* const input, T* output, int pad_top, int pad_bottom, int pad_left, int pad_right,
                        const int* original_size)
{
    static_assert(std::is_trivially_copyable<T>::value, "Reflect_101 moves raw pixels");
    blur_b200::check(blur_b200_reflect_101_u8_host(reinterpret_cast<const uint8_t*>(input), reinterpret_cast<uint8_t*>(output),
                                                   pad_top, pad_bottom, pad_left, pad_right, original_size,
                                                   C * static_cast<int>(sizeof(T))),
                     "Reflect_101");
}

// Utils.hpp:159-184 / :186-210.  T = interleaved element type, U = plane element type, any of
// uint8_t, uint16_t, int32_t, float; the reference's rounding rule (0.5f only when the destination is
// integral and the source is not) is applied by the library.  <uint8_t, float> is the instantiation the
// blur itself uses (Source.cpp:188,268,310,389,461,567).
namespace blur_b200 {
template <typename T> struct type_code;
template <> struct type_code<uint8_t> { static constexpr int value = BLUR_B200_T_U8; };
template <> struct type_code<uint16_t> { static constexpr int value = BLUR_B200_T_U16; };
template <> struct type_code<int32_t> { static constexpr int value = BLUR_B200_T_I32; };
template <> struct type_code<float> { static constexpr int value = BLUR_B200_T_F32; };
}  // namespace blur_b200

template <typename T, typename U>
inline void deinterleave_BGR(const T* const interleaved_BGR, U** const deinterleaved_BGR, const uint32_t total_size)
{
    blur_b200::check(blur_b200_deinterleave_typed_host(interleaved_BGR, deinterleaved_BGR[0], deinterleaved_BGR[1], deinterleaved_BGR[2],
                                                       total_size, blur_b200::type_code<T>::value, blur_b200::type_code<U>::value),
                     "deinterleave_BGR");
}
template <typename T, typename U>
inline void interleave_BGR(const U** const deinterleaved_BGR, T* const interleaved_BGR, const uint32_t total_size)
{
    blur_b200::check(blur_b200_interleave_typed_host(deinterleaved_BGR[0], deinterleaved_BGR[1], deinterleaved_BGR[2], interleaved_BGR,
                                                     total_size, blur_b200::type_code<U>::value, blur_b200::type_code<T>::value),
                     "interleave_BGR");
}

// flip_block<T,C>(in, out, w, h) as called at Source.cpp:367,384,540,562 (<float,1>): in is h x w pixels of
// C elements, out is w x h.  Any trivially copyable T with C*sizeof(T) in {1,2,3,4,6,8,12,16} bytes.
template <typename T, int C>
inline void flip_block(const T* in, T* out, int w, int h)
{
    static_assert(std::is_trivially_copyable<T>::value, "flip_block moves raw pixels");
    blur_b200::check(blur_b200_flip_block_pixels_host(in, out, w, h, C * static_cast<int>(sizeof(T))), "flip_block");
}
