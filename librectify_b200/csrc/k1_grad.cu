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

// One tap row of both filters for one output pixel column c: `first` starts the sums (tap row 0).
// Hx row a: -K[ar][0], -K[ar][1], 0, +K[ar][1], +K[ar][0];  Hy row a: sign(a-2) * {K[0][ac], K[1][ac], K[2][ac], K[1][ac], K[0][ac]}
template <int A>
__device__ __forceinline__ void tap_row(const float (&m)[3][2][8], float (&sx)[4], float (&sy)[4])
{
    constexpr int ar = A < 2 ? A : 4 - A;       // 0,1,2,1,0
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        float x = (A == 0) ? -m[ar][0][c] : __fsub_rn(sx[c], m[ar][0][c]);
        x = __fsub_rn(x, m[ar][1][c + 1]);
        x = __fadd_rn(x, m[ar][1][c + 3]);
        x = __fadd_rn(x, m[ar][0][c + 4]);
        sx[c] = x;
        if (A != 2) {
            constexpr int ac = ar;              // 0 or 1 here
            float y;
            if (A < 2) {
                y = (A == 0) ? -m[0][ac][c] : __fsub_rn(sy[c], m[0][ac][c]);
                y = __fsub_rn(y, m[1][ac][c + 1]);
                y = __fsub_rn(y, m[2][ac][c + 2]);
                y = __fsub_rn(y, m[1][ac][c + 3]);
                y = __fsub_rn(y, m[0][ac][c + 4]);
            } else {
                y = __fadd_rn(sy[c], m[0][ac][c]);
                y = __fadd_rn(y, m[1][ac][c + 1]);
                y = __fadd_rn(y, m[2][ac][c + 2]);
                y = __fadd_rn(y, m[1][ac][c + 3]);
                y = __fadd_rn(y, m[0][ac][c + 4]);
            }
            sy[c] = y;
        }
    }
}

// The row loop is deliberately NOT unrolled: a fully unrolled 12-row body is ~70 KB of SASS, far beyond the 32 KB
// instruction cache level (ncu: 24 % instruction-cache misses, issue slots 65-75 % busy).  Instead the five partially
// accumulated output rows live in "role" registers — role a holds the output row whose next tap row is a — and shift by
// one role per input row, so the loop body (~8 KB) stays cache resident.
template <int RPT, bool INTERIOR>
__device__ __forceinline__ float compute_strip(const float * tile, const Planes & P, int slot, int x0, int y0, int lane, int warp)
{
    const int ty0 = warp * RPT;                 // first output row of this thread inside the tile
    const int xq = x0 + lane * 4;               // first output column
    float rx[5][4], ry[5][4];                   // role registers
    float tmax = 0.0f;
    // column validity of the four output columns (2-px border), hoisted out of the row loop
    const bool vc0 = INTERIOR || (xq + 0 >= 2 && xq + 0 < P.w - 2), vc1 = INTERIOR || (xq + 1 >= 2 && xq + 1 < P.w - 2);
    const bool vc2 = INTERIOR || (xq + 2 >= 2 && xq + 2 < P.w - 2), vc3 = INTERIOR || (xq + 3 >= 2 && xq + 3 < P.w - 2);
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) { rx[a][c] = 0.f; ry[a][c] = 0.f; }

    float * pdx = P.dx + slot * P.stride; float * pdy = P.dy + slot * P.stride; float * pmag = P.mag + slot * P.stride;
    uint8_t * pbin = P.bin + slot * P.stride;
    const float * trow = tile + ty0 * K1_SW + lane * 4;

#pragma unroll 1
    for (int i = 0; i < RPT + 4; ++i, trow += K1_SW) {
        // staged column k <-> image column x0 - 4 + k; this thread needs image columns xq-2 .. xq+5 = staged 4*lane+2 .. +9
        const float4 q0 = *reinterpret_cast<const float4 *>(trow);
        const float4 q1 = *reinterpret_cast<const float4 *>(trow + 4);
        const float4 q2 = *reinterpret_cast<const float4 *>(trow + 8);
        const float p[8] = {q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y};
        float m[3][2][8];                       // m[a'][b'][j] = p[j] * |Hx(a',b')|
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            m[0][0][j] = __fmul_rn(p[j], LRB_K00); m[0][1][j] = __fmul_rn(p[j], LRB_K01);
            m[1][0][j] = __fmul_rn(p[j], LRB_K10); m[1][1][j] = __fmul_rn(p[j], LRB_K11);
            m[2][0][j] = __fmul_rn(p[j], LRB_K20); m[2][1][j] = __fmul_rn(p[j], LRB_K21);
        }
        // input row i is tap row a of output row i - a; rows outside [0, RPT) are skipped (warp-uniform branches)
        if (i >= 4)                 tap_row<4>(m, rx[4], ry[4]);
        if (i >= 3 && i < RPT + 3)  tap_row<3>(m, rx[3], ry[3]);
        if (i >= 2 && i < RPT + 2)  tap_row<2>(m, rx[2], ry[2]);
        if (i >= 1 && i < RPT + 1)  tap_row<1>(m, rx[1], ry[1]);
        if (i < RPT)                tap_row<0>(m, rx[0], ry[0]);
        // output row j = i - 4 is complete
        if (i >= 4) {
            const int y = y0 + ty0 + i - 4;
            if (INTERIOR || (y < P.h && xq < P.wq)) {
                const bool vy = INTERIOR || ((y >= 2) && (y < P.h - 2));
                float4 odx, ody, omag; unsigned b0, b1, b2, b3;
                finish_pixel<INTERIOR>(rx[4][0], ry[4][0], vy && vc0, odx.x, ody.x, omag.x, b0);
                finish_pixel<INTERIOR>(rx[4][1], ry[4][1], vy && vc1, odx.y, ody.y, omag.y, b1);
                finish_pixel<INTERIOR>(rx[4][2], ry[4][2], vy && vc2, odx.z, ody.z, omag.z, b2);
                finish_pixel<INTERIOR>(rx[4][3], ry[4][3], vy && vc3, odx.w, ody.w, omag.w, b3);
                const size_t o = (size_t)y * P.pitch + xq;
                *reinterpret_cast<float4 *>(pdx + o) = odx;
                *reinterpret_cast<float4 *>(pdy + o) = ody;
                *reinterpret_cast<float4 *>(pmag + o) = omag;
                *reinterpret_cast<unsigned *>(pbin + o) = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
                tmax = fmaxf(tmax, fmaxf(fmaxf(omag.x, omag.y), fmaxf(omag.z, omag.w)));
            }
        }
        // every row moves on to its next tap row
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            rx[4][c] = rx[3][c]; rx[3][c] = rx[2][c]; rx[2][c] = rx[1][c]; rx[1][c] = rx[0][c];
            ry[4][c] = ry[3][c]; ry[3][c] = ry[2][c]; ry[2][c] = ry[1][c]; ry[1][c] = ry[0][c];
        }
    }
    return tmax;
}

template <int RPT>
__global__ void __launch_bounds__(256, 2)
k_grad5x5_bin(const __grid_constant__ CUtensorMap tmap, Planes P, FrameCtrl * ctrl, int tma_x, int tma_y)
{
    constexpr int TH = 8 * RPT, SH = TH + 4;
    __shared__ __align__(128) float tile[SH * K1_SW];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned wmax[8];

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
        for (int k = 1; k < 8; ++k) v = max(v, wmax[k]);
        if (v) atomicMax(&ctrl[slot].gmax_bits, v);
    }
}

} // namespace

void launch_grad(const CUtensorMap & tmap, const Planes & P, FrameCtrl * ctrl, int n_slots, int tma_x, int tma_y, cudaStream_t st)
{
    constexpr int RPT = 8;
    dim3 grid((P.w + K1_TW - 1) / K1_TW, (P.h + 8 * RPT - 1) / (8 * RPT), n_slots);
    k_grad5x5_bin<RPT><<<grid, 256, 0, st>>>(tmap, P, ctrl, tma_x, tma_y);
}

int grad_tile_w() { return K1_SW; }
int grad_tile_h() { return 8 * 8 + 4; }

} // namespace lrb
