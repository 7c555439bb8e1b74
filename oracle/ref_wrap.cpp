/*
 * oracle/ref_wrap.cpp — TEST INFRASTRUCTURE.
 *
 * Compiles the reference's own Source.cpp + Utils.hpp UNMODIFIED, from where they lie under
 * /root/reference (path passed as -DREF_SOURCE="..."), into oracle/_ref/libblurref*.so and
 * exposes C entry points around the reference's functions.  Nothing from the reference is
 * copied into this repository; the build recipe is oracle/Makefile.
 *
 * The three absent third-party submodules and OpenCV are replaced by the API shims under
 * oracle/shims: the pipeline logic that runs here is the
 * reference's; the FFT arithmetic underneath is oracle/cpu_fft.c.
 */
#define main reference_main_unused
#include REF_SOURCE
#undef main

extern "C" {

/* Source.cpp:60-73 */
int ref_gaussian_window(double sigma, int max_width) { return gaussian_window(sigma, max_width); }

/* Source.cpp:75-102; out must hold max(width, fft_length) floats */
void ref_get_gaussian(float *out, double sigma, int width, int fft_length)
{
    std::vector<float> k(fft_length ? fft_length : width);
    getGaussian(k, sigma, width, fft_length);
    std::copy(k.begin(), k.end(), out);
}

/* Utils.hpp:141-157 */
int ref_is_valid_size(int n) { return isValidSize(n); }
int ref_nearest_transform_size(int n) { return nearestTransformSize(n); }

/* Utils.hpp:159-210 */
void ref_deinterleave_u8_f32(const uint8_t *in, float *b, float *g, float *r, uint32_t total)
{
    float *p[3] = {b, g, r};
    deinterleave_BGR(in, p, total);
}
void ref_interleave_f32_u8(const float *b, const float *g, const float *r, uint8_t *out, uint32_t total)
{
    const float *p[3] = {b, g, r};
    interleave_BGR(p, out, total);
}

/* Utils.hpp:212-243 */
void ref_reflect_101_u8(const uint8_t *in, uint8_t *out, int top, int bottom, int left, int right,
                        int rows, int cols, int channels)
{
    const int sz[2] = {rows, cols};
    if (channels == 1) Reflect_101<uint8_t, 1>(in, out, top, bottom, left, right, sz);
    else if (channels == 3) Reflect_101<uint8_t, 3>(in, out, top, bottom, left, right, sz);
    else if (channels == 4) Reflect_101<uint8_t, 4>(in, out, top, bottom, left, right, sz);
    else std::abort();
}

/* Test(image, flag, nsmooth), Source.cpp:574-608: 5 pocketfft_1D, 4 fastboxblur, 3 pffft_,
 * 2 pocketfft_2D (DFT_image variant if this TU was built with -DDFT_image). In place. */
void ref_test(uint8_t *bgr, int rows, int cols, int flag, double nsmooth)
{
    cv::Mat m;
    m.data = bgr;
    m.size[0] = rows;
    m.size[1] = cols;
    m.nch = 3;
    Test(m, flag, nsmooth);
}

int ref_has_dft_image(void)
{
#ifdef DFT_image
    return 1;
#else
    return 0;
#endif
}

/* bench.py --impl reference sets this itself: torchrun exports OMP_NUM_THREADS=1 */
void ref_set_num_threads(int n)
{
#ifdef OMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int ref_num_threads(void)
{
#ifdef OMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
}
