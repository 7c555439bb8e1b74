/*
 * oracle/shims/pocketfft/pocketfft_hdronly.h — TEST INFRASTRUCTURE.
 *
 * API stand-in for mreineck/pocketfft @ bb87ca50 (absent git submodule).  Only the two
 * entry points Source.cpp uses, float instantiation, contiguous row-major data:
 *   r2c: Source.cpp:212,229,238,332,343,361,377     c2r: Source.cpp:263,363,379
 * Published semantics restated: r2c transforms the LAST axis in `axes` real->half-complex
 * and the remaining axes complex; c2r is its inverse order; both unscaled times `fct`.
 * Implemented in ../shim_impl.cpp on top of oracle/cpu_fft.c (with a per-length plan
 * cache, as pocketfft has).
 */
#ifndef SHIM_POCKETFFT_H
#define SHIM_POCKETFFT_H
#include <complex>
#include <cstddef>
#include <vector>

namespace pocketfft {
using shape_t = std::vector<size_t>;
using stride_t = std::vector<ptrdiff_t>;
constexpr bool FORWARD = true, BACKWARD = false;

void shim_r2c_f32(const shape_t &shape_in, const stride_t &stride_in, const stride_t &stride_out,
                  const shape_t &axes, bool forward, const float *data_in,
                  std::complex<float> *data_out, float fct);
void shim_c2r_f32(const shape_t &shape_out, const stride_t &stride_in, const stride_t &stride_out,
                  const shape_t &axes, bool forward, const std::complex<float> *data_in,
                  float *data_out, float fct);

inline void r2c(const shape_t &shape_in, const stride_t &stride_in, const stride_t &stride_out,
                const shape_t &axes, bool forward, const float *data_in,
                std::complex<float> *data_out, float fct, size_t /*nthreads*/ = 1)
{
    shim_r2c_f32(shape_in, stride_in, stride_out, axes, forward, data_in, data_out, fct);
}
inline void c2r(const shape_t &shape_out, const stride_t &stride_in, const stride_t &stride_out,
                const shape_t &axes, bool forward, const std::complex<float> *data_in,
                float *data_out, float fct, size_t /*nthreads*/ = 1)
{
    shim_c2r_f32(shape_out, stride_in, stride_out, axes, forward, data_in, data_out, fct);
}
}  // namespace pocketfft
#endif
