// DRAM fetch granularity probe: how many DRAM bytes does a 32-byte-per-row strided read cost, and does
// the neighbouring 32-byte sector (same 64-byte pair) hit in L2 when it is requested later?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_sector_probe l2_sector_probe.cu
// Run under: ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ldcg(const float4* p)
{
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ldcg128(const float4* p)
{
    float4 v;
    asm volatile("ld.global.cg.L2::128B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// The buffer is rows x pitch bytes.  A "strip" = 64-byte column group; a unit = (strip, 1024-row block).
// MODE 0: read sector 0 of every strip only.
// MODE 1: read sector 0 then sector 1 of the same rows in the next instruction.
// MODE 2: read sector 0 of unit u, sector 1 of unit u one unit later (software pipelined).
// MODE 3: 64 bytes per row in one instruction (4 lanes x 16 B).
// MODE 4: MODE 2 with L2::128B on the loads.
// MODE 5: MODE 2 plus prefetch.global.L2 of the neighbour sector at the time of the first read.
template <int MODE>
__global__ void __launch_bounds__(128) probe(const char* __restrict__ buf, size_t pitch, int rows, int nstrips, float* sink)
{
    const int nbl = rows / 1024;
    const long long nunits = static_cast<long long>(nstrips) * nbl;
    const int q0 = static_cast<int>(nunits * blockIdx.x / gridDim.x), q1 = static_cast<int>(nunits * (blockIdx.x + 1) / gridDim.x);
    const int tid = threadIdx.x;
    float acc = 0.f;
    auto rd = [&](int q, int sector) {
        const int strip = q / nbl, b = q % nbl;
        const char* base = buf + static_cast<size_t>(b) * 1024 * pitch + static_cast<size_t>(strip) * 64 + sector * 32;
        if (MODE == 3) {
            const int l = tid & 3, t = tid >> 2;   // 32 rows per instruction
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                const float4 v = ldcg(reinterpret_cast<const float4*>(base + static_cast<size_t>(t + 32 * j) * pitch + 16 * l));
                acc += v.x + v.y + v.z + v.w;
            }
        } else {
            const int l = tid & 1, t = tid >> 1;   // 64 rows per instruction
#pragma unroll 8
            for (int j = 0; j < 16; ++j) {
                const float4* p = reinterpret_cast<const float4*>(base + static_cast<size_t>(t + 64 * j) * pitch + 16 * l);
                const float4 v = MODE == 4 ? ldcg128(p) : ldcg(p);
                if (MODE == 5 && sector == 0 && l == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(p) + 32));
                acc += v.x + v.y + v.z + v.w;
            }
        }
    };
    if (MODE == 0) for (int q = q0; q < q1; ++q) rd(q, 0);
    if (MODE == 1) for (int q = q0; q < q1; ++q) { rd(q, 0); rd(q, 1); }
    if (MODE == 2 || MODE == 4 || MODE == 5) {
        for (int q = q0; q <= q1; ++q) {
            if (q < q1) rd(q, 0);
            if (q > q0) rd(q - 1, 1);
        }
    }
    if (MODE == 3) for (int q = q0; q < q1; ++q) rd(q, 0);
    if (acc == 123.456f) *sink = acc;
}

int main()
{
    const int rows = 16384;
    const size_t pitch = 16384 * 4;          // one channel plane of the 16384^2 float intermediate
    const int nstrips = static_cast<int>(pitch / 64);
    char* buf; float* sink;
    cudaMalloc(&buf, rows * pitch);
    cudaMalloc(&sink, 4);
    cudaMemset(buf, 0, rows * pitch);
    const int grid = 148 * 3;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto run = [&](int mode) {
        float ms = 0;
        for (int it = 0; it < 2; ++it) {
            cudaEventRecord(e0);
            switch (mode) {
            case 0: probe<0><<<grid, 128>>>(buf, pitch, rows, nstrips, sink); break;
            case 1: probe<1><<<grid, 128>>>(buf, pitch, rows, nstrips, sink); break;
            case 2: probe<2><<<grid, 128>>>(buf, pitch, rows, nstrips, sink); break;
            case 3: probe<3><<<grid, 128>>>(buf, pitch, rows, nstrips, sink); break;
            case 4: probe<4><<<grid, 128>>>(buf, pitch, rows, nstrips, sink); break;
            case 5: probe<5><<<grid, 128>>>(buf, pitch, rows, nstrips, sink); break;
            }
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
        }
        const double useful = (mode == 0 ? 0.5 : 1.0) * rows * pitch;
        printf("mode %d: %.3f ms, %.1f GB/s useful (%s)\n", mode, ms, useful / ms * 1e-6, cudaGetErrorString(cudaGetLastError()));
    };
    for (int m = 0; m < 6; ++m) run(m);
    return 0;
}
