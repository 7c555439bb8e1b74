// fft_core.cuh — per-thread building blocks of the shared-memory FFT convolution.
//
// One "transform" is a length-N complex FFT that carries TWO real image lines (line a in the
// real part, line b in the imaginary part).  Because the Gaussian kernel spectrum is real
// (centred symmetric taps, Source.cpp:96-100,362,423) multiplying the packed spectrum by it
// filters both lines at once and no split/unpack step is ever needed.
//
// Decomposition: N = r0 * r1 * ... * r_{m-1} with r_{m-1} == 32.
//   forward  = decimation in frequency, natural order in, digit-scrambled order out;
//   multiply = by the kernel spectrum stored in that same scrambled order (the reference's
//              "unordered layout" idea, README.md "forwardToInternalLayout");
//   inverse  = decimation in time, scrambled in, natural out.
// The last forward pass, the multiply and the first inverse pass all act on the same 32
// register-resident values, so they are one fused step ("mid").
//
// Shared-memory layout: logical element p lives at float2 slot p + (p >> 5).  Every stride used
// by the strided passes is a multiple of 32, so the skew costs no index arithmetic there, and it
// makes the 32-contiguous-elements-per-thread access of the mid step bank-conflict free.
//
// All functions are host/device so tools/emulate.cpp can execute the exact same code on the CPU.
#pragma once
#include <stdint.h>
#include "dft_codelets.cuh"

#ifdef __CUDACC__
#define BLUR_DEV_INLINE __host__ __device__ __forceinline__
#else
#define BLUR_DEV_INLINE inline
#ifndef BLUR_FLOAT2_DEFINED
#define BLUR_FLOAT2_DEFINED
struct float2 { float x, y; };
static inline float2 make_float2(float a, float b) { float2 r; r.x = a; r.y = b; return r; }
#endif
#endif

namespace blurb200 {

constexpr int kMaxPass = 4;
constexpr int kLastRadix = 32;
constexpr int kMaxThreads = 576;   // CTA size cap of fft_conv_kernel: 18 warps

// Device-visible description of one FFT length.
struct FftGeom {
    int N;                  // FFT length (valid size: 32 * 2^a 3^b 5^c)
    int npass;              // number of passes, 2..kMaxPass; radix[npass-1] == 32
    int radix[kMaxPass];
    int M[kMaxPass];        // logical stride of pass i: N / (r0*...*ri); M[npass-1] == 1
    int W[kMaxPass];        // work items per transform in pass i: N / radix[i]
    uint64_t magicM[kMaxPass];  // ceil(2^40 / M[i])  (exact division for n*M < 2^40)
    uint64_t magicW[kMaxPass];  // ceil(2^40 / W[i])
    int slots;              // float2 slots per transform in shared memory: N + N/32
};

// radix-1 "pass": lets N == 32 use the same two-pass skeleton
template <bool INV> struct Dft<1, INV> { static BLUR_DEV_INLINE void run(float*, float*) {} };

BLUR_DEV_INLINE uint32_t fastdiv(uint32_t n, uint64_t magic) { return (uint32_t)(((uint64_t)n * magic) >> 40); }

BLUR_DEV_INLINE void cmul(float& ar, float& ai, float br, float bi)
{
    const float r = ar * br - ai * bi;
    const float i = ar * bi + ai * br;
    ar = r; ai = i;
}

// x[q] *= w^q for q = 1..R-1 (conj(w) if CONJ).  Two interleaved power chains (odd/even) keep the
// dependent-multiply depth at R/2 and the register cost at a handful of values.
template <int R, bool CONJ>
BLUR_DEV_INLINE void apply_twiddle_powers(float* xr, float* xi, float wr, float wi)
{
    if (R < 2) return;
    if (CONJ) wi = -wi;
    cmul(xr[1], xi[1], wr, wi);
    if (R < 3) return;
    float w2r = wr * wr - wi * wi, w2i = 2.f * wr * wi;   // w^2
    float er = w2r, ei = w2i;                              // even chain: w^2, w^4, ...
    float orr = wr, oi = wi;                               // odd chain:  w^1, w^3, ...
#pragma unroll
    for (int q = 2; q < R; ++q) {
        if (q & 1) {
            cmul(orr, oi, w2r, w2i);
            cmul(xr[q], xi[q], orr, oi);
        } else {
            if (q > 2) cmul(er, ei, w2r, w2i);
            cmul(xr[q], xi[q], er, ei);
        }
    }
}

// ---- strided passes on shared memory -----------------------------------------------------
// `s` points at the transform's slot 0.  u in [0, N/R).  tw[k] = exp(-2*pi*i*k/N).

// IO functor contract (first forward pass reads through it, last inverse pass writes):
//   Ctx  begin(int unit) const;                                       once per work item
//   void load (const Ctx&, int n, float& re, float& im) const;        n in [0, N)
//   void store(const Ctx&, int n, float re, float im) const;

template <int R, bool FIRST, class IO>
BLUR_DEV_INLINE void fwd_pass(float2* __restrict__ s, const FftGeom& g, int pass, int u, const float2* __restrict__ tw,
                              const IO& io, int unit)
{
    const int M = g.M[pass];
    const int blk = FIRST ? 0 : (int)fastdiv((uint32_t)u, g.magicM[pass]);
    const int t = u - blk * M;
    const int base_log = blk * R * M + t;
    const int base = base_log + (base_log >> 5);
    const int Mp = M + (M >> 5);
    float xr[R], xi[R];
    const typename IO::Ctx ctx = FIRST ? io.begin(unit) : typename IO::Ctx();
#pragma unroll
    for (int j = 0; j < R; ++j) {
        if (FIRST) {
            io.load(ctx, t + j * M, xr[j], xi[j]);
        } else {
            const float2 v = s[base + j * Mp];
            xr[j] = v.x; xi[j] = v.y;
        }
    }
    Dft<R, false>::run(xr, xi);
    if (R > 1) {
        const float2 w = tw[t * (g.N / (R * M))];
        apply_twiddle_powers<R, false>(xr, xi, w.x, w.y);
    }
#pragma unroll
    for (int q = 0; q < R; ++q) s[base + q * Mp] = make_float2(xr[q], xi[q]);
}

template <int R, bool LAST, class IO>
BLUR_DEV_INLINE void inv_pass(float2* __restrict__ s, const FftGeom& g, int pass, int u, const float2* __restrict__ tw,
                              const IO& io, int unit)
{
    const int M = g.M[pass];
    const int blk = LAST ? 0 : (int)fastdiv((uint32_t)u, g.magicM[pass]);
    const int t = u - blk * M;
    const int base_log = blk * R * M + t;
    const int base = base_log + (base_log >> 5);
    const int Mp = M + (M >> 5);
    float xr[R], xi[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
        const float2 v = s[base + j * Mp];
        xr[j] = v.x; xi[j] = v.y;
    }
    if (R > 1) {
        const float2 w = tw[t * (g.N / (R * M))];
        apply_twiddle_powers<R, true>(xr, xi, w.x, w.y);
    }
    Dft<R, true>::run(xr, xi);
    const typename IO::Ctx ctx = LAST ? io.begin(unit) : typename IO::Ctx();
#pragma unroll
    for (int q = 0; q < R; ++q) {
        if (LAST) io.store(ctx, t + q * M, xr[q], xi[q]);
        else s[base + q * Mp] = make_float2(xr[q], xi[q]);
    }
}

// Fused last-forward / multiply / first-inverse step.  u in [0, N/32).  kt is the kernel
// spectrum (already scaled by 1/N) in scrambled order, transposed: kt[q * (N/32) + u].
BLUR_DEV_INLINE void mid_pass(float2* __restrict__ s, const FftGeom& g, int u, const float* __restrict__ kt)
{
    constexpr int R = kLastRadix;
    const int W = g.N / R;
    const int base = u * (R + 1);
    float xr[R], xi[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
        const float2 v = s[base + j];
        xr[j] = v.x; xi[j] = v.y;
    }
    Dft<R, false>::run(xr, xi);
#pragma unroll
    for (int q = 0; q < R; ++q) {
        const float k = kt[q * W + u];
        xr[q] *= k; xi[q] *= k;
    }
    Dft<R, true>::run(xr, xi);
#pragma unroll
    for (int j = 0; j < R; ++j) s[base + j] = make_float2(xr[j], xi[j]);
}

// Frequency index held at logical position p after the forward passes:
// p = q0*M0 + q1*M1 + ... + q_{m-1}  ->  k = q0 + r0*(q1 + r1*(q2 + ...)).
inline int scrambled_freq(const FftGeom& g, int p)
{
    int k = 0, mul = 1;
    for (int i = 0; i < g.npass; ++i) {
        const int q = p / g.M[i];
        p -= q * g.M[i];
        k += q * mul;
        mul *= g.radix[i];
    }
    return k;
}

}  // namespace blurb200
