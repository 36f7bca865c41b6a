// pencil.cuh — exact float32 helpers of the line-pencil model shared by the grouping kernels (k6_ransac.cu, k_hough.cu):
// cross product / hypothesis (reference src/line_pencil.cpp:101-108), inclination / error (src/geometry.cpp:214-229,
// src/line_pencil.cpp:131-134) and a block-wide flag scan.  Every operation is an explicitly rounded intrinsic in the
// reference's evaluation order.
#pragma once
#include "common.cuh"

namespace lrb {

namespace {

// ---- small exact helpers ----------------------------------------------------------------------------------------------
__device__ __forceinline__ float sum3(float a, float b, float c) { return __fadd_rn(a, __fadd_rn(b, c)); }   // fixed-size-3 reduction order

struct Hyp { float p0, p1, p2, px, py; bool ideal; };

__device__ __forceinline__ Hyp make_hyp(float p0, float p1, float p2)
{
    Hyp h; h.p0 = p0; h.p1 = p1; h.p2 = p2;
    h.ideal = fabsf(p2) < kEps;                                                 // geometry.cpp:219
    h.px = h.ideal ? p0 : __fdiv_rn(p0, p2);
    h.py = h.ideal ? p1 : __fdiv_rn(p1, p2);
    return h;
}

// 1 - |cos| of the angle between the segment direction and the direction anchor -> vanishing point
__device__ __forceinline__ float line_error(const Hyp & h, float ax, float ay, float dx, float dy)
{
    float vx, vy;
    if (h.ideal) { vx = h.px; vy = h.py; }
    else { vx = __fadd_rn(-ax, h.px); vy = __fadd_rn(-ay, h.py); }
    const float z = __fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy));
    if (z > 0.f) { const float nr = __fsqrt_rn(z); vx = __fdiv_rn(vx, nr); vy = __fdiv_rn(vy, nr); }
    const float inc = fabsf(__fadd_rn(__fmul_rn(vx, dx), __fmul_rn(vy, dy)));
    return __fadd_rn(-inc, 1.0f);
}

__device__ __forceinline__ void cross3(const float * A, const float * B, float & p0, float & p1, float & p2)
{
    p0 = __fsub_rn(__fmul_rn(A[1], B[2]), __fmul_rn(A[2], B[1]));
    p1 = __fsub_rn(__fmul_rn(A[2], B[0]), __fmul_rn(A[0], B[2]));
    p2 = __fsub_rn(__fmul_rn(A[0], B[1]), __fmul_rn(A[1], B[0]));
}

// block-wide exclusive scan of 0/1 flags (blockDim = 1024); returns position, total through `total`
__device__ __forceinline__ int block_scan_flag(int flag, int * wsum, int & total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    const int inw = __popc(m & ((1u << lane) - 1));
    if (lane == 0) wsum[warp] = __popc(m);
    __syncthreads();
    if (warp == 0) {
        int v = lane < nw ? wsum[lane] : 0;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, s); if (lane >= s) v += u; }
        wsum[lane] = v;
    }
    __syncthreads();
    const int pos = inw + (warp ? wsum[warp - 1] : 0);
    total = wsum[nw - 1];
    __syncthreads();
    return pos;
}

} // namespace

} // namespace lrb
