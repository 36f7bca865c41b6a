// k1_grad.cu — K1 `grad5x5_bin`: fused image gradients, magnitude, orientation bin and frame maximum.
//
// Replaces, for one frame per blockIdx.z:
//   conv_2d x2 + image_gradients          reference src/filter.cpp:81-98, src/line_detector.cpp:41-49
//   first two loops of gradient_directions reference src/line_detector.cpp:140-156
//   max(mag) of find_line_segments        reference src/line_detector.cpp:209
//
// Arithmetic contract (bit-exact against the reference compiled without FMA contraction / vectorisation):
//   dx(r,c) = sum over the 5x5 window in row-major tap order of img*Hx, each product and each sum rounded separately;
//   the window is a direct 25-tap correlation (NOT separable: separable evaluation rounds differently).
//   The taps satisfy Hx(a,b) = -Hx(a,4-b) = Hx(4-a,b), Hx(a,2) = 0 and Hy = Hx^T exactly, so only six distinct
//   magnitudes exist.  Each input pixel is multiplied once by those six magnitudes and the rounded products are
//   shared by every output and by both filters (IEEE multiplication is sign-symmetric, so x*(-k) == -(x*k)).
//   Zero taps contribute x*0 = +-0, which can only change the sign of an exactly-zero sum when the input holds
//   negative zeros; they are skipped ("K1 exactness").
//
// Data movement: a (4+128+4) x (8*RPT+4) halo tile is staged into shared memory by ONE cp.async.bulk.tensor (TMA) per
// CTA; the copy engine zero-fills outside the image.  Each thread owns a 4-column x RPT-row strip, slides down the
// tile keeping five partially accumulated output rows in registers, and writes dx/dy/mag as 128-bit and bin as
// 32-bit coalesced stores.  HBM traffic is the compulsory 4 B read + 13 B written per pixel.
#include <cstdlib>
#include <cstdio>
#include "common.cuh"
#include "tma.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int K1_TW = 128;          // tile width  (32 lanes x 4 columns)
constexpr int K1_HALO_L = 4;        // staged columns left of the tile: 2 needed, 4 keeps the TMA box origin 16-byte aligned
constexpr int K1_SW = K1_HALO_L + K1_TW + 4;   // staged width (544 B rows: multiple of 16 B as TMA requires)

template <bool INTERIOR>
__device__ __forceinline__ void finish_pixel(float dx, float dy, bool valid, float & odx, float & ody, float & omag, unsigned & obin)
{
    if (!INTERIOR) { dx = valid ? dx : 0.0f; dy = valid ? dy : 0.0f; }        // conv_2d leaves the 2-px border at 0 (filter.cpp:81-98)
    odx = dx; ody = dy;
    omag = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));      // line_detector.cpp:47
    // eight orientation energies |dx*sin(t_i) + dy*cos(t_i)|, t_i = float(i*pi)/8 (line_detector.cpp:142-146).
    // sin/cos pairs share magnitudes: c2 = s2, c6 = -s2, c7 = -c1, s0 = 0, c0 = 1, s4 = 1.
    const float S1 = 0x1.87de2cp-2f, C1 = 0x1.d906bcp-1f, S2 = 0x1.6a09e6p-1f, S3 = 0x1.d906bcp-1f, C3 = 0x1.87de2ap-2f;
    const float C4 = -0x1.777a5cp-25f, S5 = 0x1.d906bep-1f, C5 = -0x1.87de28p-2f, S7 = 0x1.87de2ep-2f;
    const float dxs2 = __fmul_rn(dx, S2), dys2 = __fmul_rn(dy, S2), dyc1 = __fmul_rn(dy, C1);
    float g[8];
    g[0] = fabsf(dy);                                                        // dx*0 + dy*1
    g[1] = fabsf(__fadd_rn(__fmul_rn(dx, S1), dyc1));
    g[2] = fabsf(__fadd_rn(dxs2, dys2));
    g[3] = fabsf(__fadd_rn(__fmul_rn(dx, S3), __fmul_rn(dy, C3)));
    g[4] = fabsf(__fadd_rn(dx, __fmul_rn(dy, C4)));                          // dx*1 + dy*c4
    g[5] = fabsf(__fadd_rn(__fmul_rn(dx, S5), __fmul_rn(dy, C5)));
    g[6] = fabsf(__fsub_rn(dxs2, dys2));                                     // dy*(-s2) == -(dy*s2)
    g[7] = fabsf(__fsub_rn(__fmul_rn(dx, S7), dyc1));                        // dy*(-c1) == -(dy*c1)
    // first strict maximum with the running maximum initialised to 0 (line_detector.cpp:152-156)
    float gm = g[0]; unsigned b = 0;                                         // g[0] > 0 selects bin 0, which is also the default
#pragma unroll
    for (int i = 1; i < 8; ++i) { b = (g[i] > gm) ? (unsigned)i : b; gm = fmaxf(gm, g[i]); }
    obin = b;
}

// ---- packed pairs -------------------------------------------------------------------------------------------------
// Blackwell's add.rn.f32x2 (SASS FADD2) adds two independent, individually rounded float pairs per instruction: the
// result is bit-identical to two scalar adds and costs one issue slot instead of two (the FP32-pipe cycles are the
// same: scripts/ubench/fp32_pipes.cu).  Products stay scalar: ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into
// FFMA2 even under --fmad=false, which would round differently; a scalar FMUL feeding a packed add is never contracted.
typedef unsigned long long f2;
__device__ __forceinline__ f2 pk(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(f2 v, float & lo, float & hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 sub2(f2 a, f2 b) { f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// finish_pixel for two neighbouring pixels at once: identical arithmetic, the eight sums per pixel as packed adds
template <bool INTERIOR>
__device__ __forceinline__ void finish_pair(f2 dxp, f2 dyp, bool v0, bool v1, float & odx0, float & odx1, float & ody0, float & ody1,
                                            float & omag0, float & omag1, unsigned & ob0, unsigned & ob1)
{
    float dx0, dx1, dy0, dy1;
    upk(dxp, dx0, dx1); upk(dyp, dy0, dy1);
    if (!INTERIOR) {                                                         // conv_2d leaves the 2-px border at 0 (filter.cpp:81-98)
        dx0 = v0 ? dx0 : 0.0f; dy0 = v0 ? dy0 : 0.0f; dx1 = v1 ? dx1 : 0.0f; dy1 = v1 ? dy1 : 0.0f;
        dxp = pk(dx0, dx1); dyp = pk(dy0, dy1);
    }
    odx0 = dx0; odx1 = dx1; ody0 = dy0; ody1 = dy1;
    float s0, s1;
    upk(add2(pk(__fmul_rn(dx0, dx0), __fmul_rn(dx1, dx1)), pk(__fmul_rn(dy0, dy0), __fmul_rn(dy1, dy1))), s0, s1);
    omag0 = __fsqrt_rn(s0); omag1 = __fsqrt_rn(s1);                          // line_detector.cpp:47
    // the eight orientation energies (see finish_pixel for the constants and the shared products)
    const float S1 = 0x1.87de2cp-2f, C1 = 0x1.d906bcp-1f, S2 = 0x1.6a09e6p-1f, S3 = 0x1.d906bcp-1f, C3 = 0x1.87de2ap-2f;
    const float C4 = -0x1.777a5cp-25f, S5 = 0x1.d906bep-1f, C5 = -0x1.87de28p-2f, S7 = 0x1.87de2ep-2f;
#define LRB_P2(a, k) pk(__fmul_rn(a##0, k), __fmul_rn(a##1, k))
    const f2 dxs2 = LRB_P2(dx, S2), dys2 = LRB_P2(dy, S2), dyc1 = LRB_P2(dy, C1);
    float g0[8], g1[8];
    g0[0] = fabsf(dy0); g1[0] = fabsf(dy1);                                  // dx*0 + dy*1
    upk(add2(LRB_P2(dx, S1), dyc1), g0[1], g1[1]);
    upk(add2(dxs2, dys2), g0[2], g1[2]);
    upk(add2(LRB_P2(dx, S3), LRB_P2(dy, C3)), g0[3], g1[3]);
    upk(add2(dxp, LRB_P2(dy, C4)), g0[4], g1[4]);                            // dx*1 + dy*c4
    upk(add2(LRB_P2(dx, S5), LRB_P2(dy, C5)), g0[5], g1[5]);
    upk(sub2(dxs2, dys2), g0[6], g1[6]);                                     // dy*(-s2) == -(dy*s2)
    upk(sub2(LRB_P2(dx, S7), dyc1), g0[7], g1[7]);                           // dy*(-c1) == -(dy*c1)
#undef LRB_P2
    // first strict maximum with the running maximum initialised to 0 (line_detector.cpp:152-156)
    float m0 = g0[0], m1 = g1[0]; unsigned b0 = 0, b1 = 0;
#pragma unroll
    for (int i = 1; i < 8; ++i) {
        const float a0 = fabsf(g0[i]), a1 = fabsf(g1[i]);
        b0 = (a0 > m0) ? (unsigned)i : b0; m0 = fmaxf(m0, a0);
        b1 = (a1 > m1) ? (unsigned)i : b1; m1 = fmaxf(m1, a1);
    }
    ob0 = b0; ob1 = b1;
}

// The products of one input row, as the column pairs the two output-column pairs (q = 0: columns 0,1; q = 1: columns
// 2,3) need.  e* start at an even staged column j = 0,2,4,6 (index j/2), o* at an odd one j = 1,3,5 (index (j-1)/2).
// kXY = p * |Hx(X,Y)|; Hx row a = {-k[a'][0], -k[a'][1], 0, +k[a'][1], +k[a'][0]}, a' = min(a, 4-a); Hy = Hx^T.
struct RowProducts {
    f2 e00[4], e10[4], e20[4], e01[4], e21[4];      // e21: only [1], [2] are used
    f2 o10[3], o01[3], o11[3], o21[3];
};

__device__ __forceinline__ void row_products(const float (&p)[8], RowProducts & R)
{
    float k00[8], k10[8], k20[8], k01[8], k11[8], k21[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        k00[j] = __fmul_rn(p[j], LRB_K00); k10[j] = __fmul_rn(p[j], LRB_K10); k20[j] = __fmul_rn(p[j], LRB_K20);
        k01[j] = __fmul_rn(p[j], LRB_K01); k11[j] = __fmul_rn(p[j], LRB_K11); k21[j] = __fmul_rn(p[j], LRB_K21);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        R.e00[i] = pk(k00[2 * i], k00[2 * i + 1]); R.e10[i] = pk(k10[2 * i], k10[2 * i + 1]);
        R.e20[i] = pk(k20[2 * i], k20[2 * i + 1]); R.e01[i] = pk(k01[2 * i], k01[2 * i + 1]);
        R.e21[i] = pk(k21[2 * i], k21[2 * i + 1]);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        R.o10[i] = pk(k10[2 * i + 1], k10[2 * i + 2]); R.o01[i] = pk(k01[2 * i + 1], k01[2 * i + 2]);
        R.o11[i] = pk(k11[2 * i + 1], k11[2 * i + 2]); R.o21[i] = pk(k21[2 * i + 1], k21[2 * i + 2]);
    }
}

// Tap row A of Hx for output-column pair q, continuing the partial sum `acc` (ignored for A = 0).
// Reference order per output: ((((acc - t0) - t1) + t3) + t4); for A = 0, (-t0 - t1) == -(t0 + t1) exactly.
template <int A>
__device__ __forceinline__ f2 hx_row(const RowProducts & R, int q, f2 acc)
{
    constexpr int ar = A < 2 ? A : 4 - A;
    const f2 * e = ar == 0 ? R.e00 : (ar == 1 ? R.e10 : R.e20);
    const f2 * o = ar == 0 ? R.o01 : (ar == 1 ? R.o11 : R.o21);
    f2 x;
    if (A == 0) { x = add2(e[q], o[q]); x = sub2(o[q + 1], x); }
    else        { x = sub2(acc, e[q]); x = sub2(x, o[q]); x = add2(x, o[q + 1]); }
    return add2(x, e[q + 2]);
}

// Tap row A of Hy (A != 2) for pair q.  Rows 0 and 1 subtract all their taps: the running value is kept NEGATED
// (-(a + b) == (-a) - b exactly), rows 3 and 4 add; row 3 turns the sign with t0 - acc == (-acc) + t0.
template <int A>
__device__ __forceinline__ f2 hy_row(const RowProducts & R, int q, f2 acc)
{
    constexpr int ac = A < 2 ? A : 4 - A;
    const f2 * e0 = ac == 0 ? R.e00 : R.e01;
    const f2 * o1 = ac == 0 ? R.o10 : R.o11;
    const f2 * e2 = ac == 0 ? R.e20 : R.e21;
    f2 y;
    if (A == 0)      y = add2(e0[q], o1[q]);
    else if (A == 3) { y = sub2(e0[q], acc); y = add2(y, o1[q]); }
    else             { y = add2(acc, e0[q]); y = add2(y, o1[q]); }
    y = add2(y, e2[q + 1]);
    y = add2(y, o1[q + 1]);
    return add2(y, e0[q + 2]);
}

// One thread: 4 output columns (two pairs) x RPT output rows.  The loop over input rows is rolled (a fully unrolled body
// is far beyond the instruction cache); the five partially accumulated output rows live in "role" registers — role a
// holds the output row whose next tap row is a+1 — and every tap row writes its result into the NEXT role, evaluated
// from role 4 down to role 0, so the shift costs no moves.
template <int RPT, bool INTERIOR>
__device__ __forceinline__ float compute_strip(const float * tile, const Planes & P, int slot, int x0, int y0, int lane, int warp)
{
    const int ty0 = warp * RPT;                 // first output row of this thread inside the tile
    const int xq = x0 + lane * 4;               // first output column
    f2 rx[4][2], ry[4][2];                      // roles 0..3 (after tap rows 0..3) for the two column pairs; ry[2] only waits
    float tmax = 0.0f;
    // column validity of the four output columns (2-px border), hoisted out of the row loop
    const bool vc0 = INTERIOR || (xq + 0 >= 2 && xq + 0 < P.w - 2), vc1 = INTERIOR || (xq + 1 >= 2 && xq + 1 < P.w - 2);
    const bool vc2 = INTERIOR || (xq + 2 >= 2 && xq + 2 < P.w - 2), vc3 = INTERIOR || (xq + 3 >= 2 && xq + 3 < P.w - 2);
#pragma unroll
    for (int a = 0; a < 4; ++a) { rx[a][0] = rx[a][1] = 0ull; ry[a][0] = ry[a][1] = 0ull; }

    float2 * pdxy = P.dxy + slot * P.stride; float * pmag = P.mag + slot * P.stride;
    uint8_t * pbin = P.bin + slot * P.stride;
    const float * trow = tile + ty0 * K1_SW + lane * 4;

#pragma unroll 1
    for (int i = 0; i < RPT + 4; ++i, trow += K1_SW) {
        // staged column k <-> image column x0 - 4 + k; this thread needs image columns xq-2 .. xq+5 = staged 4*lane+2 .. +9
        const float4 q0 = *reinterpret_cast<const float4 *>(trow);
        const float4 q1 = *reinterpret_cast<const float4 *>(trow + 4);
        const float4 q2 = *reinterpret_cast<const float4 *>(trow + 8);
        const float p[8] = {q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y};
        RowProducts R;
        row_products(p, R);
        // input row i is tap row a of output row i - a: role 3 -> finished row, role 2 -> 3, ..., fresh row -> role 0
        // (tap row a is needed for 0 <= i - a < RPT; the guards are warp-uniform and skip the work for rows outside the strip)
        f2 fx[2] = {0ull, 0ull}, fy[2] = {0ull, 0ull};
        if (i >= 4) {
#pragma unroll
            for (int q = 0; q < 2; ++q) { fx[q] = hx_row<4>(R, q, rx[3][q]); fy[q] = hy_row<4>(R, q, ry[3][q]); }
        }
        if (i >= 3 && i < RPT + 3) {
#pragma unroll
            for (int q = 0; q < 2; ++q) { rx[3][q] = hx_row<3>(R, q, rx[2][q]); ry[3][q] = hy_row<3>(R, q, ry[2][q]); }
        }
        if (i >= 2 && i < RPT + 2) {
#pragma unroll
            for (int q = 0; q < 2; ++q) { rx[2][q] = hx_row<2>(R, q, rx[1][q]); ry[2][q] = ry[1][q]; }   // Hy's middle tap row is zero
        }
        if (i >= 1 && i < RPT + 1) {
#pragma unroll
            for (int q = 0; q < 2; ++q) { rx[1][q] = hx_row<1>(R, q, rx[0][q]); ry[1][q] = hy_row<1>(R, q, ry[0][q]); }
        }
        if (i < RPT) {
#pragma unroll
            for (int q = 0; q < 2; ++q) { rx[0][q] = hx_row<0>(R, q, 0ull); ry[0][q] = hy_row<0>(R, q, 0ull); }
        }
        // output row j = i - 4 is complete
        if (i >= 4) {
            const int y = y0 + ty0 + i - 4;
            if (INTERIOR || (y < P.h && xq < P.wq)) {
                const bool vy = INTERIOR || ((y >= 2) && (y < P.h - 2));
                float4 odx, ody, omag; unsigned b0, b1, b2, b3;
                finish_pair<INTERIOR>(fx[0], fy[0], vy && vc0, vy && vc1, odx.x, odx.y, ody.x, ody.y, omag.x, omag.y, b0, b1);
                finish_pair<INTERIOR>(fx[1], fy[1], vy && vc2, vy && vc3, odx.z, odx.w, ody.z, ody.w, omag.z, omag.w, b2, b3);
                const size_t o = (size_t)y * P.pitch + xq;
                *reinterpret_cast<float4 *>(pdxy + o) = make_float4(odx.x, ody.x, odx.y, ody.y);
                *reinterpret_cast<float4 *>(pdxy + o + 2) = make_float4(odx.z, ody.z, odx.w, ody.w);
                *reinterpret_cast<float4 *>(pmag + o) = omag;
                *reinterpret_cast<unsigned *>(pbin + o) = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
                tmax = fmaxf(tmax, fmaxf(fmaxf(omag.x, omag.y), fmaxf(omag.z, omag.w)));
            }
        }
    }
    return tmax;
}

template <int RPT, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB)
k_grad5x5_bin(const __grid_constant__ CUtensorMap tmap, Planes P, FrameCtrl * ctrl, int tma_x, int tma_y)
{
    constexpr int TH = WARPS * RPT, SH = TH + 4;
    __shared__ __align__(128) float tile[SH * K1_SW];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned wmax[WARPS];

    const int slot = blockIdx.z;
    const int x0 = blockIdx.x * K1_TW, y0 = blockIdx.y * TH;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // TMA requires the box origin to be 16-byte aligned along x (an unaligned start raises "illegal instruction" on
    // B200), so the staged tile starts 4 columns (not 2) left of the first output column; negative / out-of-range
    // coordinates are legal and zero-filled, which is the reference's border rule.
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, SH * K1_SW * sizeof(float));
        tma_load_3d(tile, &tmap, x0 - K1_HALO_L + tma_x, y0 - 2 + tma_y, slot, &bar);
    }
    mbar_wait(&bar, 0);

    // tiles that touch neither the 2-px border nor the right / bottom edge skip all validity logic
    const bool interior = (x0 >= 2) && (x0 + K1_TW <= P.w - 2) && (y0 >= 2) && (y0 + TH <= P.h - 2);
    const float tmax = interior ? compute_strip<RPT, true>(tile, P, slot, x0, y0, lane, warp)
                                : compute_strip<RPT, false>(tile, P, slot, x0, y0, lane, warp);

    // frame maximum of mag: mag >= 0, so the uint order of the bit patterns is the float order
    unsigned um = __float_as_uint(tmax);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) um = max(um, __shfl_xor_sync(0xffffffffu, um, s));
    if (lane == 0) wmax[warp] = um;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned v = wmax[0];
#pragma unroll
        for (int k = 1; k < WARPS; ++k) v = max(v, wmax[k]);
        if (v) atomicMax(&ctrl[slot].gmax_bits, v);
    }
}

// Persistent form: a resident grid walks the tiles of all frames with two TMA stages per CTA, so the next tile is in
// flight while the current one is evaluated (the one-shot form spends 17 % of its stall samples in the mbarrier wait at CTA
// start).  8 warps x 8 rows per 128 x 64 tile, two CTAs per SM: 2 x 2 x 37 KB of shared memory, inside the 66 % split that
// keeps the L1 the flood kernels of the other engines need (see launch_nms in k2_nms.cu).
template <int RPT, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB)
k_grad5x5_bin_persistent(const __grid_constant__ CUtensorMap tmap, Planes P, FrameCtrl * ctrl, int tma_x, int tma_y,
                         int tiles_x, int tiles_y, int n_tiles)
{
    constexpr int TH = WARPS * RPT, SH = TH + 4, STAGE = SH * K1_SW * (int)sizeof(float);
    static_assert(STAGE % 128 == 0, "TMA destinations must stay 128-byte aligned");
    extern __shared__ unsigned char k1_smem_raw[];
    unsigned char * sm = k1_smem_raw + ((128u - (smem_u32(k1_smem_raw) & 127u)) & 127u);
    uint64_t * bar = reinterpret_cast<uint64_t *>(sm + 2 * STAGE);
    unsigned * wmax = reinterpret_cast<unsigned *>(bar + 2);                 // [2][WARPS]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int per_frame = tiles_x * tiles_y;
    auto issue = [&](int t, int stage) {        // thread 0 only
        const int slot = t / per_frame, r = t - slot * per_frame, by = r / tiles_x, bx = r - by * tiles_x;
        mbar_arrive_expect_tx(&bar[stage], STAGE);
        tma_load_3d(sm + stage * STAGE, &tmap, bx * K1_TW - K1_HALO_L + tma_x, by * TH - 2 + tma_y, slot, &bar[stage]);
    };
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_fence_init();
        if ((int)blockIdx.x < n_tiles) issue(blockIdx.x, 0);
    }
    __syncthreads();
    int it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int stage = it & 1;
        if (threadIdx.x == 0 && t + (int)gridDim.x < n_tiles) issue(t + gridDim.x, stage ^ 1);   // stage^1 was released by the barrier that ended iteration it-1
        const int slot = t / per_frame, r = t - slot * per_frame, by = r / tiles_x, bx = r - by * tiles_x;
        const int x0 = bx * K1_TW, y0 = by * TH;
        const float * tile = reinterpret_cast<const float *>(sm + stage * STAGE);
        mbar_wait(&bar[stage], (it >> 1) & 1);
        const bool interior = (x0 >= 2) && (x0 + K1_TW <= P.w - 2) && (y0 >= 2) && (y0 + TH <= P.h - 2);
        const float tmax = interior ? compute_strip<RPT, true>(tile, P, slot, x0, y0, lane, warp)
                                    : compute_strip<RPT, false>(tile, P, slot, x0, y0, lane, warp);
        unsigned um = __float_as_uint(tmax);
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) um = max(um, __shfl_xor_sync(0xffffffffu, um, s));
        if (lane == 0) wmax[stage * WARPS + warp] = um;
        __syncthreads();                        // everyone is done with this stage; the warp maxima are complete
        if (threadIdx.x == 0) {
            unsigned v = wmax[stage * WARPS];
#pragma unroll
            for (int k = 1; k < WARPS; ++k) v = max(v, wmax[stage * WARPS + k]);
            if (v) atomicMax(&ctrl[slot].gmax_bits, v);
        }
    }
}

} // namespace

void launch_grad(const CUtensorMap & tmap, const Planes & P, FrameCtrl * ctrl, int n_slots, int tma_x, int tma_y, cudaStream_t st)
{
    // both shapes cover a 128 x 64 tile (the TMA box of grad_tile_w() x grad_tile_h()): 8 warps x 8 rows or 4 warps x 16 rows
    // per thread; the taller strip amortises the 4-row halo (products of 20 input rows per 16 output rows instead of 12 per 8)
    static const int cfg = [] { const char * v = getenv("LRB_K1_CFG"); return v ? atoi(v) : 1; }();
    dim3 grid((P.w + K1_TW - 1) / K1_TW, (P.h + 63) / 64, n_slots);
    if (cfg >= 10) {                            // LRB_K1_CFG=10: persistent, double-buffered — the documented alternative (profiles/r1_sweeps.md)
        auto kern = k_grad5x5_bin_persistent<8, 8, 2>;
        constexpr int smem = 2 * (64 + 4) * K1_SW * (int)sizeof(float) + 64 + 8 * 8 + 128;
        static PerDeviceOnce once; static int n_sm[PerDeviceOnce::kMaxDev];
        const int d = once.run([&] {
            const int dv = PerDeviceOnce::dev();
            cudaError_t e1 = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaError_t e2 = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 66);
            if (e1 != cudaSuccess || e2 != cudaSuccess) fprintf(stderr, "[lrb] K1: shared-memory opt-in failed on device %d\n", dv);
            n_sm[dv] = 148; cudaDeviceGetAttribute(&n_sm[dv], cudaDevAttrMultiProcessorCount, dv);
        });
        const int n_tiles = (int)(grid.x * grid.y * grid.z);
        const int blocks = n_tiles < 2 * n_sm[d] ? n_tiles : 2 * n_sm[d];
        kern<<<blocks, 256, smem, st>>>(tmap, P, ctrl, tma_x, tma_y, (int)grid.x, (int)grid.y, n_tiles);
        return;
    }
    LRB_CARVEOUT((k_grad5x5_bin<16, 4, 4>));
    k_grad5x5_bin<16, 4, 4><<<grid, 128, 0, st>>>(tmap, P, ctrl, tma_x, tma_y);
}

int grad_tile_w() { return K1_SW; }
int grad_tile_h() { return 8 * 8 + 4; }

} // namespace lrb
