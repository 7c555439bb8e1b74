// fp_pipes.cu — instruction-throughput microbenchmark for the fp32 pipes of sm_100a:
// scalar FADD / FMUL / FFMA versus the packed FADD2 / FMUL2 / FFMA2 (PTX add/mul/fma.f32x2).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp_pipes fp_pipes.cu ; run on a B200.
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 2048
#define CHAINS 8

template <int OP> __global__ void __launch_bounds__(256) kern(float* out, float s0, float s1)
{
    float a[CHAINS], b[CHAINS];
    unsigned long long A[CHAINS];
    const float c0 = s0, c1 = s1;
    unsigned long long C0, C1;
    asm("mov.b64 %0, {%1, %2};" : "=l"(C0) : "f"(s0), "f"(s0));
    asm("mov.b64 %0, {%1, %2};" : "=l"(C1) : "f"(s1), "f"(s1));
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
        a[i] = threadIdx.x * 1e-3f + i; b[i] = a[i] + 0.5f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(A[i]) : "f"(a[i]), "f"(b[i]));
    }
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) {
            if (OP == 0) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c0));
            if (OP == 1) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c0));
            if (OP == 2) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c0), "f"(c1));          // 3 distinct regs
            if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(a[i]) : "f"(c0));                    // 2 distinct regs
            if (OP == 4) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(A[i]) : "l"(C0));
            if (OP == 5) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(A[i]) : "l"(C0));
            if (OP == 6) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(A[i]) : "l"(C0), "l"(C1));
            if (OP == 7) asm volatile("fma.rn.f32 %0, %0, 0f3F800001, %1;" : "+f"(a[i]) : "f"(c1));           // immediate multiplier
            if (OP == 8) {  // mixed: FADD2 + 2 scalar FADD (alternating), models a radix-4 butterfly mix
                asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(A[i]) : "l"(C0));
                asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c0));
            }
            if (OP == 9) {  // FFMA2 + MOV-ish ALU op (lop3) to see co-issue with the alu pipe
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(A[i]) : "l"(C0), "l"(C1));
                asm volatile("xor.b32 %0, %0, 0x80000000;" : "+r"(*(unsigned*)&b[i]));
            }
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(A[i]));
        acc += a[i] + b[i] + lo + hi;
    }
    if (acc == 123.456f) out[0] = acc;
}

template <int OP> void run(const char* name, int instr_per_slot, int flops_per_instr)
{
    int dev = 0; cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    float* d; cudaMalloc(&d, 4);
    const int blocks = p.multiProcessorCount * 8;   // 8 x 256 threads = 64 warps per SM
    kern<OP><<<blocks, 256>>>(d, 1.0000001f, 1e-9f);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<OP><<<blocks, 256>>>(d, 1.0000001f, 1e-9f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double winstr = (double)blocks * 8 /*warps*/ * ITER * CHAINS * instr_per_slot;
    const double per_sm_per_ns = winstr / p.multiProcessorCount / (ms * 1e6);
    printf("%-34s %8.3f ms  %6.3f warp-instr/ns/SM  (= %5.2f /clk @1.965GHz)  %7.1f TFLOP/s\n", name, ms, per_sm_per_ns,
           per_sm_per_ns / 1.965, winstr * 32 * flops_per_instr / (ms * 1e-3) / 1e12 / instr_per_slot * (instr_per_slot == 2 ? 1.0 : 1.0));
    cudaFree(d);
}

int main()
{
    run<0>("FADD  r,r", 1, 1);
    run<1>("FMUL  r,r", 1, 1);
    run<2>("FFMA  r,r,r (3 distinct)", 1, 2);
    run<3>("FFMA  r,r,r (2 distinct)", 1, 2);
    run<7>("FFMA  r,imm,r", 1, 2);
    run<4>("FADD2", 1, 2);
    run<5>("FMUL2", 1, 2);
    run<6>("FFMA2 (3 distinct)", 1, 4);
    run<8>("FADD2 + FADD pair", 2, 1);
    run<9>("FFMA2 + LOP3 pair", 2, 2);
    return 0;
}
