// cplx2.cuh — complex arithmetic on (re, im) pairs held in one 64-bit register pair.
//
// sm_100a has packed fp32 instructions (PTX add/mul/fma.rn.f32x2 -> SASS FADD2 / FMUL2 / FFMA2)
// whose operands accept a LO_HI swap, per-half negation and a broadcast scalar.  ptxas folds the
// make_float2(...) shuffles below into those modifiers, so every function here is ONE packed
// instruction (cmul / cmulc: two).  Measured on B200 (tools/microbench/fp_pipes.cu): a packed
// instruction occupies the FMA pipe for two cycles, i.e. the flop rate equals the scalar one, but
// it needs half the issue slots — which is what lets the shared-memory traffic and address
// arithmetic of the FFT passes issue underneath the butterflies.
//
// Host versions (CPU emulator in tests/emul) use fmaf so both sides round identically.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define BLUR_HD __host__ __device__ __forceinline__
#else
#include <cmath>
#define BLUR_HD inline
#ifndef BLUR_FLOAT2_DEFINED
#define BLUR_FLOAT2_DEFINED
struct float2 { float x, y; };
static inline float2 make_float2(float a, float b) { float2 r; r.x = a; r.y = b; return r; }
#endif
#endif

namespace blurb200 {

// store one complex value as a single 64-bit access.  (A plain `*p = v` makes ptxas copy the
// register pair before every STS.64.)
BLUR_HD void cstore(float2* p, float2 v)
{
#ifdef __CUDA_ARCH__
    unsigned long long u;
    asm("mov.b64 %0, {%1, %2};" : "=l"(u) : "f"(v.x), "f"(v.y));
    asm volatile("st.shared.b64 [%0], %1;" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(p))), "l"(u) : "memory");
#else
    *p = v;
#endif
}

// a + b
BLUR_HD float2 cadd(float2 a, float2 b)
{
#ifdef __CUDA_ARCH__
    return __fadd2_rn(a, b);
#else
    return make_float2(a.x + b.x, a.y + b.y);
#endif
}
// a - b
BLUR_HD float2 csub(float2 a, float2 b)
{
#ifdef __CUDA_ARCH__
    return __fadd2_rn(a, make_float2(-b.x, -b.y));
#else
    return make_float2(a.x - b.x, a.y - b.y);
#endif
}
// a + i*b
BLUR_HD float2 caddi(float2 a, float2 b)
{
#ifdef __CUDA_ARCH__
    return __fadd2_rn(a, make_float2(-b.y, b.x));
#else
    return make_float2(a.x - b.y, a.y + b.x);
#endif
}
// a - i*b
BLUR_HD float2 csubi(float2 a, float2 b)
{
#ifdef __CUDA_ARCH__
    return __fadd2_rn(a, make_float2(b.y, -b.x));
#else
    return make_float2(a.x + b.y, a.y - b.x);
#endif
}
// c * a   (c real)
BLUR_HD float2 cscale(float2 a, float c)
{
#ifdef __CUDA_ARCH__
    return __fmul2_rn(a, make_float2(c, c));
#else
    return make_float2(a.x * c, a.y * c);
#endif
}
// a + c*b   (c real)
BLUR_HD float2 cfmar(float2 a, float c, float2 b)
{
#ifdef __CUDA_ARCH__
    return __ffma2_rn(b, make_float2(c, c), a);
#else
    return make_float2(fmaf(b.x, c, a.x), fmaf(b.y, c, a.y));
#endif
}
// a + i*c*b   (c real)
BLUR_HD float2 cfmai(float2 a, float c, float2 b)
{
#ifdef __CUDA_ARCH__
    return __ffma2_rn(make_float2(b.y, b.x), make_float2(-c, c), a);
#else
    return make_float2(fmaf(b.y, -c, a.x), fmaf(b.x, c, a.y));
#endif
}
// a * w
BLUR_HD float2 cmul(float2 a, float2 w)
{
#ifdef __CUDA_ARCH__
    const float2 t = __fmul2_rn(a, make_float2(w.x, w.x));
    return __ffma2_rn(make_float2(a.y, a.x), make_float2(-w.y, w.y), t);
#else
    return make_float2(fmaf(a.y, -w.y, a.x * w.x), fmaf(a.x, w.y, a.y * w.x));
#endif
}
// a * conj(w)
BLUR_HD float2 cmulc(float2 a, float2 w)
{
#ifdef __CUDA_ARCH__
    const float2 t = __fmul2_rn(a, make_float2(w.x, w.x));
    return __ffma2_rn(make_float2(a.y, a.x), make_float2(w.y, -w.y), t);
#else
    return make_float2(fmaf(a.y, w.y, a.x * w.x), fmaf(a.x, -w.y, a.y * w.x));
#endif
}

}  // namespace blurb200
