/*
 * oracle/blur_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, float32, optional OpenMP) of the hot path of
 * michelerenzullo/Blur_algorithms @ 2f873a4e.  Every function cites the
 * reference lines it follows.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product path (blur_algorithms_b200/) never does.
 *
 * Parity status:
 *   - sizing / taps / reflect / (de)interleave: pinned against the reference's
 *     verbatim Utils.hpp and Source.cpp:60-102 compiled into oracle/_ref.
 *   - Gaussian pipelines: pinned against Source.cpp compiled UNMODIFIED into
 *     oracle/_ref against API shims for the absent pffft/pocketfft/FastBoxBlur
 *     submodules (pipeline logic is the reference's; FFT arithmetic is ours).
 *   - DFT_image: pinned against the reference's 23 golden PNG pairs.
 *   - fastboxblur: PARITY UNPINNED (source absent from the tree) — declared
 *     spec, SURVEY.md §8c / Appendix A.7.
 */
#ifndef BLUR_ORACLE_H
#define BLUR_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Source.cpp:60-73 */
int orc_gaussian_window(double sigma, int max_width);
/* Source.cpp:75-102; kernel must hold max(width, fft_length) floats */
void orc_get_gaussian(float *kernel, double sigma, int width, int fft_length);
/* Utils.hpp:141-148, 150-157 */
int orc_is_valid_size(int n);
int orc_nearest_transform_size(int n);
/* Utils.hpp:159-184, 186-210 (uint8 <-> float instantiation) */
void orc_deinterleave_u8_f32(const uint8_t *in, float *b, float *g, float *r, uint32_t total);
void orc_interleave_f32_u8(const float *b, const float *g, const float *r, uint8_t *out, uint32_t total);
/* Utils.hpp:212-243, uint8 with C channels */
void orc_reflect_101_u8(const uint8_t *in, uint8_t *out, int top, int bottom, int left, int right,
                        int rows, int cols, int channels);
/* flip_block<float,1>: call sites Source.cpp:367,384,540,562 (in: h x w, out: w x h) */
void orc_flip_block_f32(const float *in, float *out, int w, int h);

/* Derived sizes, Source.cpp:434-457.  out[0]=kSize out[1]=pad out[2]=sizes[0] (col FFT
 * length) out[3]=sizes[1] (row FFT length) out[4..5]=trailing zeros per axis */
void orc_gauss_sizes(int rows, int cols, double sigma, int out[6]);

/* `#define boxblur` build of the reference (Source.cpp:12) as a run-time switch: the Gaussian pipelines
 * below take the reference's boxblur branches (tent kernel = box convolved with itself, kLen =
 * (int)(nsmooth'^2), pad = (kLen-1)/2*2).  orc_box_kernel_1d: Source.cpp:128-140. */
void orc_set_boxblur(int on);
void orc_box_kernel_1d(float *kernel, int kLen, int fft_length);

/* nyquist_mode: 0 = exact multiply on all bins (pocketfft_1D, Source.cpp:280-392),
 *               1 = pffft ordered-layout quirk (pffft_, Source.cpp:414-427, 429-570).
 * In place on interleaved 3-channel uint8.  If rowpass_out != NULL it receives the
 * float32 result of the row pass, planar [3][rows][cols] (the `resf` of Source.cpp:536). */
int orc_gauss_blur_1d(uint8_t *bgr, int rows, int cols, double sigma, int nyquist_mode,
                      float *rowpass_out);
/* same, but also returns the float32 planes just before interleave_BGR (planar [3][rows][cols]) */
int orc_gauss_blur_1d_f32(const uint8_t *bgr, int rows, int cols, double sigma, int nyquist_mode,
                          float *rowpass_out, float *final_out);

/* pocketfft_2D, Source.cpp:143-277 (blur branch).  In place. */
int orc_gauss_blur_2d(uint8_t *bgr, int rows, int cols, double sigma);

/* DFT_image branch, Source.cpp:240-252 (+ 268-276).  nsmooth decides the padding exactly as
 * pocketfft_2D does; spec_f32 (optional) receives the float spectrum of the PADDED planes,
 * planar [3][sizes0][sizes1]; out_u8 receives the cropped interleaved uint8 image
 * (negatives saturate to 0 as the reference fixtures show).  padded_sizes[2] returned. */
int orc_dft_image(const uint8_t *bgr, uint8_t *out_u8, int rows, int cols, double nsmooth,
                  float *spec_f32, int padded_sizes[2]);

/* fastboxblur(in, w, h, channels, ksize, passes): call site Source.cpp:587.
 * DECLARED SPEC (upstream source absent): r = (ksize-1)/2, width k = 2r+1, reflect-101
 * borders, horizontal then vertical, uint8 stored after each 1-D pass with
 * round-half-up (2*acc + k) / (2k), repeated `passes` times, in place.
 * Returns nonzero if r > min(w,h)-1. */
int orc_fastboxblur(uint8_t *img, int w, int h, int channels, int ksize, int passes);

int orc_num_threads(void);
void orc_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
