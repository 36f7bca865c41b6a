// Microbenchmark: issue rate of scalar FADD/FMUL versus packed FADD2 (add.rn.f32x2) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -o fp32_pipes fp32_pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2(u64 a, u64 b){ u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
template <int MODE>
__global__ void __launch_bounds__(256) k(float * out, float s, int iters)
{
    float a[16]; u64 p[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = ((u64)__float_as_uint(a[2 * i + 1]) << 32) | __float_as_uint(a[2 * i]);
    const u64 s2 = ((u64)__float_as_uint(s) << 32) | __float_as_uint(s);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {            // 16 scalar FADD
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = __fadd_rn(a[i], s);
        } else if (MODE == 1) {     // 8 FADD2 (same flops)
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = add2(p[i], s2);
        } else if (MODE == 2) {     // 16 scalar FMUL
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = __fmul_rn(a[i], s);
        } else if (MODE == 3) {     // 8 FADD + 8 FMNMX (alu pipe) interleaved
#pragma unroll
            for (int i = 0; i < 8; ++i) { a[i] = __fadd_rn(a[i], s); a[8 + i] = fmaxf(a[8 + i], a[i]); }
        } else if (MODE == 4) {     // 16 FMNMX
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaxf(a[i], a[(i + 1) & 15] + 0.f);
        }
    }
    float r = 0.f;
    if (MODE == 1) { for (int i = 0; i < 8; ++i) r += __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32)); }
    else for (int i = 0; i < 16; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char * name, float * out, double ops_per_iter)
{
    const int iters = 20000, blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(out, 1.0001f, 100);
    cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(out, 1.0001f, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warp_instr = (double)blocks * 8 * iters * ops_per_iter;
    // per SMSP per clock at 1.965 GHz (nominal; printed for orientation only)
    printf("%-28s %8.3f ms  %.3f warp-instr/clk/SMSP (at 1965 MHz)\n", name, ms, warp_instr / (148.0 * 4) / (ms * 1e-3 * 1.965e9));
}
int main()
{
    float * out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("16 FADD", out, 16);
    run<1>("8 FADD2", out, 8);
    run<2>("16 FMUL", out, 16);
    run<3>("8 FADD + 8 FMNMX", out, 16);
    run<4>("16 FMNMX(+FADD)", out, 32);
    return 0;
}
