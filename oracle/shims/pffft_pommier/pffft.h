/*
 * oracle/shims/pffft_pommier/pffft.h — TEST INFRASTRUCTURE.
 *
 * API stand-in for jpommier/pffft @ fbc40586 (absent git submodule of the reference,
 * /root/reference/.SUBMODULES.json).  Declares only the entry points Source.cpp calls
 * (Source.cpp:477-478, 485, 499, 531-533, 553-555, 565-566) with pffft's published
 * semantics: real transforms, unscaled, "ordered" layout
 *   [F0.re, F(N/2).re, F1.re, F1.im, F2.re, F2.im, ...].
 * Implemented in ../shim_impl.cpp on top of oracle/cpu_fft.c.
 *
 * The standard headers below are what the reference's translation unit gets
 * transitively from the real third-party headers / the author's libc++.
 */
#ifndef SHIM_PFFFT_H
#define SHIM_PFFFT_H
#include <algorithm>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <limits>
#include <memory>
#include <type_traits>
#include <vector>

/* libstdc++'s std::vector copy-assignment (Source.cpp:503, `kerf_1D_col = kerf_1D_row`)
 * compares allocators with != even for stateless allocators; the author's libc++ does not.
 * Utils.hpp's PFAlloc defines no comparison, so supply the (always-equal) one here. */
template <class T> class PFAlloc;
template <class T, class U> inline bool operator==(const PFAlloc<T> &, const PFAlloc<U> &) { return true; }
template <class T, class U> inline bool operator!=(const PFAlloc<T> &, const PFAlloc<U> &) { return false; }

extern "C" {
typedef struct PFFFT_Setup PFFFT_Setup;
typedef enum { PFFFT_FORWARD, PFFFT_BACKWARD } pffft_direction_t;
typedef enum { PFFFT_REAL, PFFFT_COMPLEX } pffft_transform_t;
PFFFT_Setup *pffft_new_setup(int N, pffft_transform_t transform);
void pffft_destroy_setup(PFFFT_Setup *);
void pffft_transform_ordered(PFFFT_Setup *setup, const float *input, float *output, float *work,
                             pffft_direction_t direction);
}
#endif
