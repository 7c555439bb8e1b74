// cplx2.cuh — complex arithmetic on (re, im) pairs held in one 64-bit register pair.
//
// sm_100a has packed fp32 instructions (PTX add/mul/fma.rn.f32x2 -> SASS FADD2 / FMUL2 / FFMA2)
// whose operands accept a LO_HI swap, per-half negation and a broadcast scalar.  ptxas folds the
// unpack / neg / repack sequences below into those modifiers, so every function here is ONE
// packed instruction (cmul / cmulc: two).  Measured on B200 (tools/microbench/fp_pipes.cu): a
// packed instruction occupies the FMA pipe for two cycles, i.e. the flop rate equals the scalar
// one, but it needs half the issue slots — which is what lets the shared-memory traffic and
// address arithmetic of the FFT passes issue underneath the butterflies.
//
// Under nvcc a complex value is an opaque 64-bit integer (`cplx`) from the load to the store: the
// CUDA float2 intrinsics (__fadd2_rn ...) unpack every result into two 32-bit registers, and
// ptxas then copies pairs around (two IMAD.MOV on the FMA pipe before most STS.64).  Keeping the
// value in a .b64 register end to end leaves nothing to re-pair.
//
// Host versions (CPU emulator in tests/emul, compiled by g++) use a float2 struct and fmaf so
// both sides round identically.
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

#ifdef __CUDACC__
// ------------------------------------------------------------------ device representation
typedef unsigned long long cplx;

#ifdef __CUDA_ARCH__
#define BLUR_ASM(...) asm(__VA_ARGS__)
#else
// host pass of nvcc: never executed (the kernels are the only callers), must only compile
#define BLUR_ASM(...) d = 0
#endif

BLUR_HD cplx cpack(float re, float im)
{
    cplx d;
    BLUR_ASM("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(re), "f"(im));
    return d;
}
BLUR_HD cplx czero() { return 0ull; }
BLUR_HD float cre(cplx a)
{
#ifdef __CUDA_ARCH__
    float x, y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a));
    return x;
#else
    return static_cast<float>(a);
#endif
}
BLUR_HD float cim(cplx a)
{
#ifdef __CUDA_ARCH__
    float x, y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a));
    return y;
#else
    return static_cast<float>(a);
#endif
}
// a + b
BLUR_HD cplx cadd(cplx a, cplx b)
{
    cplx d;
    BLUR_ASM("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// a - b
BLUR_HD cplx csub(cplx a, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y;\n.reg .b64 n;\nmov.b64 {x, y}, %2;\nneg.f32 x, x;\nneg.f32 y, y;\nmov.b64 n, {x, y};\n"
             "add.rn.f32x2 %0, %1, n;\n}" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// a + i*b
BLUR_HD cplx caddi(cplx a, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y;\n.reg .b64 n;\nmov.b64 {x, y}, %2;\nneg.f32 y, y;\nmov.b64 n, {y, x};\n"
             "add.rn.f32x2 %0, %1, n;\n}" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// a - i*b
BLUR_HD cplx csubi(cplx a, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y;\n.reg .b64 n;\nmov.b64 {x, y}, %2;\nneg.f32 x, x;\nmov.b64 n, {y, x};\n"
             "add.rn.f32x2 %0, %1, n;\n}" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// c * a   (c real)
BLUR_HD cplx cscale(cplx a, float c)
{
    cplx d;
    BLUR_ASM("{\n.reg .b64 cc;\nmov.b64 cc, {%2, %2};\nmul.rn.f32x2 %0, %1, cc;\n}" : "=l"(d) : "l"(a), "f"(c));
    return d;
}
// a + c*b   (c real)
BLUR_HD cplx cfmar(cplx a, float c, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .b64 cc;\nmov.b64 cc, {%2, %2};\nfma.rn.f32x2 %0, %3, cc, %1;\n}" : "=l"(d) : "l"(a), "f"(c), "l"(b));
    return d;
}
// a + i*c*b   (c real)  = a + {-b.y, b.x} * {c, c}
BLUR_HD cplx cfmai(cplx a, float c, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y;\n.reg .b64 cc, s;\nmov.b64 {x, y}, %3;\nneg.f32 y, y;\nmov.b64 s, {y, x};\n"
             "mov.b64 cc, {%2, %2};\nfma.rn.f32x2 %0, s, cc, %1;\n}" : "=l"(d) : "l"(a), "f"(c), "l"(b));
    return d;
}
// a - c*b   (c real)
BLUR_HD cplx cfmsr(cplx a, float c, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y;\n.reg .b64 cc, n;\nmov.b64 {x, y}, %3;\nneg.f32 x, x;\nneg.f32 y, y;\nmov.b64 n, {x, y};\n"
             "mov.b64 cc, {%2, %2};\nfma.rn.f32x2 %0, n, cc, %1;\n}" : "=l"(d) : "l"(a), "f"(c), "l"(b));
    return d;
}
// a - i*c*b   (c real)  = a + {b.y, -b.x} * {c, c}
BLUR_HD cplx cfmsi(cplx a, float c, cplx b)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y;\n.reg .b64 cc, s;\nmov.b64 {x, y}, %3;\nneg.f32 x, x;\nmov.b64 s, {y, x};\n"
             "mov.b64 cc, {%2, %2};\nfma.rn.f32x2 %0, s, cc, %1;\n}" : "=l"(d) : "l"(a), "f"(c), "l"(b));
    return d;
}
// 2a - s
BLUR_HD cplx c2amb(cplx a, cplx s)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 x, y, two;\n.reg .b64 cc, n;\nmov.b64 {x, y}, %2;\nneg.f32 x, x;\nneg.f32 y, y;\nmov.b64 n, {x, y};\n"
             "mov.f32 two, 0f40000000;\nmov.b64 cc, {two, two};\nfma.rn.f32x2 %0, %1, cc, n;\n}" : "=l"(d) : "l"(a), "l"(s));
    return d;
}
// a * w  = a * {wx, wx} + {-ay, ax} * {wy, wy}
BLUR_HD cplx cmul(cplx a, cplx w)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 ax, ay, wx, wy;\n.reg .b64 t, s, u, v;\nmov.b64 {ax, ay}, %1;\nmov.b64 {wx, wy}, %2;\n"
             "mov.b64 u, {wx, wx};\nmul.rn.f32x2 t, %1, u;\nneg.f32 ay, ay;\nmov.b64 s, {ay, ax};\nmov.b64 v, {wy, wy};\n"
             "fma.rn.f32x2 %0, s, v, t;\n}" : "=l"(d) : "l"(a), "l"(w));
    return d;
}
// a * conj(w)  = a * {wx, wx} + {ay, -ax} * {wy, wy}
BLUR_HD cplx cmulc(cplx a, cplx w)
{
    cplx d;
    BLUR_ASM("{\n.reg .f32 ax, ay, wx, wy;\n.reg .b64 t, s, u, v;\nmov.b64 {ax, ay}, %1;\nmov.b64 {wx, wy}, %2;\n"
             "mov.b64 u, {wx, wx};\nmul.rn.f32x2 t, %1, u;\nneg.f32 ax, ax;\nmov.b64 s, {ay, ax};\nmov.b64 v, {wy, wy};\n"
             "fma.rn.f32x2 %0, s, v, t;\n}" : "=l"(d) : "l"(a), "l"(w));
    return d;
}
#undef BLUR_ASM

#else
// ------------------------------------------------------------------ host representation (emulator)
typedef float2 cplx;

BLUR_HD cplx cpack(float re, float im) { return make_float2(re, im); }
BLUR_HD cplx czero() { return make_float2(0.f, 0.f); }
BLUR_HD float cre(cplx a) { return a.x; }
BLUR_HD float cim(cplx a) { return a.y; }
BLUR_HD cplx cadd(cplx a, cplx b) { return make_float2(a.x + b.x, a.y + b.y); }
BLUR_HD cplx csub(cplx a, cplx b) { return make_float2(a.x - b.x, a.y - b.y); }
BLUR_HD cplx caddi(cplx a, cplx b) { return make_float2(a.x - b.y, a.y + b.x); }
BLUR_HD cplx csubi(cplx a, cplx b) { return make_float2(a.x + b.y, a.y - b.x); }
BLUR_HD cplx cscale(cplx a, float c) { return make_float2(a.x * c, a.y * c); }
BLUR_HD cplx cfmar(cplx a, float c, cplx b) { return make_float2(fmaf(b.x, c, a.x), fmaf(b.y, c, a.y)); }
BLUR_HD cplx cfmai(cplx a, float c, cplx b) { return make_float2(fmaf(b.y, -c, a.x), fmaf(b.x, c, a.y)); }
BLUR_HD cplx cfmsr(cplx a, float c, cplx b) { return make_float2(fmaf(-b.x, c, a.x), fmaf(-b.y, c, a.y)); }
BLUR_HD cplx cfmsi(cplx a, float c, cplx b) { return make_float2(fmaf(b.y, c, a.x), fmaf(-b.x, c, a.y)); }
BLUR_HD cplx c2amb(cplx a, cplx s) { return make_float2(fmaf(a.x, 2.f, -s.x), fmaf(a.y, 2.f, -s.y)); }
BLUR_HD cplx cmul(cplx a, cplx w) { return make_float2(fmaf(a.y, -w.y, a.x * w.x), fmaf(a.x, w.y, a.y * w.x)); }
BLUR_HD cplx cmulc(cplx a, cplx w) { return make_float2(fmaf(a.y, w.y, a.x * w.x), fmaf(a.x, -w.y, a.y * w.x)); }
#endif

}  // namespace blurb200
