/*
 * oracle/shims/FastBoxBlur/fast_box_blur.h — TEST INFRASTRUCTURE.
 *
 * API stand-in for michelerenzullo/FastBoxBlur @ f6597c9e (absent git submodule).
 * flip_block: call sites Source.cpp:367,384,540,562.  fastboxblur: Source.cpp:587.
 * fastboxblur follows the DECLARED spec of SURVEY.md Appendix A.7 (parity unpinned:
 * upstream source is not in the tree); it forwards to the C oracle.
 */
#ifndef SHIM_FAST_BOX_BLUR_H
#define SHIM_FAST_BOX_BLUR_H
#include <cstdint>
#include <cstring>

extern "C" int orc_fastboxblur(uint8_t *img, int w, int h, int channels, int ksize, int passes);

/* in: h rows x w cols of C-channel pixels; out: w rows x h cols */
template <typename T, int C>
void flip_block(const T *in, T *out, const int w, const int h)
{
    constexpr int B = 32;
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (int y0 = 0; y0 < h; y0 += B)
        for (int x0 = 0; x0 < w; x0 += B) {
            const int y1 = y0 + B < h ? y0 + B : h, x1 = x0 + B < w ? x0 + B : w;
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x)
                    for (int c = 0; c < C; ++c)
                        out[((size_t)x * h + y) * C + c] = in[((size_t)y * w + x) * C + c];
        }
}

inline void fastboxblur(uint8_t *in, int w, int h, int channels, int ksize, int passes)
{
    orc_fastboxblur(in, w, h, channels, ksize, passes);
}
#endif
