// ubench.cu — the FP32-pipe denominators of the scoring roofline, measured on the device the benchmark runs on.
//
// K8 (k_score) is FP32-pipe bound (inner dimension 3: no tensor cores) and, like every kernel of this library, may not
// contract a multiply and an add into an FMA (bit-exact parity with the reference's strict-IEEE arithmetic).  Two peaks are
// therefore reported: FFMA (2 flop per instruction: the conventional "FP32 peak") and alternating FMUL / FADD (1 flop per
// instruction: the ceiling for FMA-free code).  16 independent chains per thread, 8 CTAs of 256 threads per SM.
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

template <int MODE>
__global__ void __launch_bounds__(256) k_ubench_fp32(float * out, float s, float t, int iters)
{
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = __fmaf_rn(a[i], s, t);
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = (i & 1) ? __fmul_rn(a[i], s) : __fadd_rn(a[i], t);
        }
    }
    float r = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

} // namespace

void launch_ubench_fp32(int mode, float * out, int blocks, int iters, cudaStream_t st)
{
    if (mode == 0) k_ubench_fp32<0><<<blocks, 256, 0, st>>>(out, 1.0001f, 0.0001f, iters);
    else           k_ubench_fp32<1><<<blocks, 256, 0, st>>>(out, 1.0001f, 0.0001f, iters);
}

} // namespace lrb
