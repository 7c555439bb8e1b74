// fft_kernels.cuh — the fused per-line FFT convolution kernel (sm_100a).
//
// One CTA owns `batch` transforms at a time in shared memory (each = two image lines packed
// as one complex sequence) and runs, per transform:
//     global load (+reflect-101 pad, zeros)  -> radix-r0 DIF butterflies -> smem
//     [radix-r1 ... strided DIF passes on smem]
//     radix-32 DIF | x kernel spectrum (scrambled order, 1/N folded in) | radix-32 DIT   (registers)
//     [strided DIT passes on smem]
//     radix-r0 DIT butterflies -> crop -> global store in the layout the next stage wants
// i.e. Source.cpp:525-536 (row pass) or :549-558 (+ :567) (column pass) without the three
// per-line heap copies, the planar float deinterleave pass, and both flip_block passes.
//
// The grid is persistent: gridDim.x = #SMs * resident CTAs, each CTA strides over the units.
#pragma once
#include "blur_io.cuh"
#include "fft_dispatch.cuh"

namespace blurb200 {


template <class IO>
__global__ void __launch_bounds__(kMaxThreads, 1)
fft_conv_kernel(const __grid_constant__ FftGeom g, const __grid_constant__ IO io,
                const float2* __restrict__ tw, const float* __restrict__ kt, const int batch)
{
    extern __shared__ float2 smem[];
    const int T = blockDim.x, tid = threadIdx.x;
    const int slots = g.slots;
    const int last = g.npass - 1;

    for (int unit0 = blockIdx.x * batch; unit0 < io.nunits; unit0 += gridDim.x * batch) {
        const int nb = min(batch, io.nunits - unit0);

        // ---- first forward pass: global -> registers -> shared
        {
            const int W = g.W[0], items = nb * W;
#define BLUR_F(R)                                                                      \
    for (int it = tid; it < items; it += T) {                                          \
        const int bt = (int)fastdiv((uint32_t)it, g.magicW[0]);                        \
        fwd_pass<R, true>(smem + bt * slots, g, 0, it - bt * W, tw, io, unit0 + bt);   \
    }
            BLUR_RADIX_SWITCH(g.radix[0], BLUR_F)
#undef BLUR_F
        }
        __syncthreads();

        // ---- remaining strided forward passes
        for (int p = 1; p < last; ++p) {
            const int W = g.W[p], items = nb * W;
#define BLUR_F(R)                                                                      \
    for (int it = tid; it < items; it += T) {                                          \
        const int bt = (int)fastdiv((uint32_t)it, g.magicW[p]);                        \
        fwd_pass<R, false>(smem + bt * slots, g, p, it - bt * W, tw, io, unit0 + bt);  \
    }
            BLUR_RADIX_SWITCH(g.radix[p], BLUR_F)
#undef BLUR_F
            __syncthreads();
        }

        // ---- radix-32 forward, multiply by the kernel spectrum, radix-32 inverse
        {
            const int W = g.W[last], items = nb * W;
            for (int it = tid; it < items; it += T) {
                const int bt = (int)fastdiv((uint32_t)it, g.magicW[last]);
                mid_pass(smem + bt * slots, g, it - bt * W, kt);
            }
        }
        __syncthreads();

        // ---- strided inverse passes
        for (int p = last - 1; p >= 1; --p) {
            const int W = g.W[p], items = nb * W;
#define BLUR_F(R)                                                                      \
    for (int it = tid; it < items; it += T) {                                          \
        const int bt = (int)fastdiv((uint32_t)it, g.magicW[p]);                        \
        inv_pass<R, false>(smem + bt * slots, g, p, it - bt * W, tw, io, unit0 + bt);  \
    }
            BLUR_RADIX_SWITCH(g.radix[p], BLUR_F)
#undef BLUR_F
            __syncthreads();
        }

        // ---- last inverse pass: shared -> registers -> global
        {
            const int W = g.W[0], items = nb * W;
#define BLUR_F(R)                                                                      \
    for (int it = tid; it < items; it += T) {                                          \
        const int bt = (int)fastdiv((uint32_t)it, g.magicW[0]);                        \
        inv_pass<R, true>(smem + bt * slots, g, 0, it - bt * W, tw, io, unit0 + bt);   \
    }
            BLUR_RADIX_SWITCH(g.radix[0], BLUR_F)
#undef BLUR_F
        }
        __syncthreads();   // smem is reused by the next unit's first pass
    }
}

}  // namespace blurb200
