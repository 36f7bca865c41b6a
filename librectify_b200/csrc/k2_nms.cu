// k2_nms.cu — K2 `nms5x5_seeds` (+ the dilated orientation mask) and K3, the exact seed ordering.
//
// Replaces, per frame:
//   maximum_filter + find_peaks           reference src/filter.cpp:29-43,161-195 (threshold src/line_detector.cpp:209)
//   binary_dilate of the 8 bin masks      reference src/filter.cpp:46-62, src/line_detector.cpp:163-171
//   std::sort of the peaks + seed_bin     reference src/filter.cpp:188, src/line_detector.cpp:216-220
//
// K2 reads mag (4 B/px) and bin (1 B/px) once, through TMA halo tiles, and writes one byte per pixel:
//   mask8(p) bit b  <=>  p is not on the 1-px border and some 3x3 neighbour has bin == b
// which is all that the reference's eight dilated, masked energy planes (66 MB at 1080p) are ever used for:
//   G_b(p) = mask8(p).b ? |dx(p) sin_b + dy(p) cos_b| : 0.
// Peaks (mag == 5x5 max, mag > max(mag)*(1-0.95f), 2-px border excluded) are appended to a per-frame list with one
// warp-aggregated atomic per warp and histogrammed into kSeedBuckets monotone magnitude buckets.
//
// K3 turns the unordered list into the reference's processing order — magnitude descending, raster index ascending
// inside exact ties (std::sort is unstable there; the tests accept any order inside a tie group) — without a general
// sort: bucket scan -> scatter -> exact rank inside the (tiny) bucket by counting.  The rank is the seed's flood
// priority and its slot in every per-seed array.
#include <cstdlib>
#include <cstdio>
#include "common.cuh"
#include "tma.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int K2_TW = 128, K2_TH = 32;
// TMA box origins must be 16-byte aligned along x: the mag tile starts 4 floats, the bin tile 16 bytes left of the tile
constexpr int K2_MW = 4 + K2_TW + 4, K2_MH = K2_TH + 4;    // mag tile: columns x0-4 .. x0+131, rows y0-2 .. y0+33
constexpr int K2_BW = 160, K2_BH = K2_TH + 2;              // bin tile: columns x0-16 .. x0+143, rows y0-1 .. y0+32

// monotone (non-decreasing in mag) bucket index, then flipped so that bucket 0 holds the largest magnitudes
__device__ __forceinline__ int seed_bucket(float mag, float thr, float gmax)
{
    const float scale = __fdiv_rn((float)kSeedBuckets, __fsub_rn(gmax, thr));
    int b = (int)__fmul_rn(__fsub_rn(mag, thr), scale);
    b = min(max(b, 0), kSeedBuckets - 1);
    return kSeedBuckets - 1 - b;
}

// same bucket with the scale factor hoisted (identical arithmetic)
__device__ __forceinline__ int seed_bucket_scaled(float mag, float thr, float scale)
{
    int b = (int)__fmul_rn(__fsub_rn(mag, thr), scale);
    b = min(max(b, 0), kSeedBuckets - 1);
    return kSeedBuckets - 1 - b;
}

__device__ __forceinline__ float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }   // one FMNMX3

// A thread owns 4 columns x K2_RPT rows of the 128 x 32 tile.  Instruction budget per pixel (SASS): ~2 shared loads,
// ~4 three-input maxima, 2 for the peak decision, ~4 for the packed-byte dilation: ~24 in all.
// One tile, staged in tmag / tbin: writes mask8, returns the peak bitmap of this thread (bit j*4+c = row j, column c).
template <int K2_RPT>
__device__ __forceinline__ unsigned nms_tile(const float * tmag, const uint8_t * tbin, const Planes & P, int slot, int x0, int y0, float tup,
                                             int lane, int warp)
{
    const int ty0 = warp * K2_RPT;              // K2_RPT output rows per thread
    const int xq = x0 + lane * 4;               // 4 output columns per thread
    // ---- 5x5 maximum, separable (max is exact and order-free) ----------------------------------------------------
    // A peak needs mag == window max and mag > thr; tup is folded into the window max, so the test is one equality.
    // The 2-px image border and everything outside the image have mag == 0 (K1 writes the reference's zero border, TMA
    // zero-fills), and 0 < tup, so no position test is needed.
    float hm[K2_RPT + 4][4];
    float4 ctr[K2_RPT];
#pragma unroll
    for (int i = 0; i < K2_RPT + 4; ++i) {
        const float * row = &tmag[(ty0 + i) * K2_MW + lane * 4];
        const float4 q0 = *reinterpret_cast<const float4 *>(row);
        const float4 q1 = *reinterpret_cast<const float4 *>(row + 4);
        const float4 q2 = *reinterpret_cast<const float4 *>(row + 8);
        // image columns xq-2 .. xq+5 = q0.z q0.w q1.x q1.y q1.z q1.w q2.x q2.y
        const float core = fmax3(q1.x, q1.y, q1.z);
        hm[i][0] = fmax3(core, q0.z, q0.w);
        hm[i][1] = fmax3(core, q0.w, q1.w);
        hm[i][2] = fmax3(core, q1.w, q2.x);
        hm[i][3] = fmax3(fmax3(q1.y, q1.z, q1.w), q2.x, q2.y);
        if (i >= 2 && i < K2_RPT + 2) ctr[i - 2] = q1;
    }
    unsigned peakmask = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        float pr[K2_RPT / 2 + 2];               // pr[k] = max(rows 2k, 2k+1, tup)
#pragma unroll
        for (int k = 0; k < K2_RPT / 2 + 2; ++k) pr[k] = fmax3(hm[2 * k][c], hm[2 * k + 1][c], tup);
#pragma unroll
        for (int k = 0; k < K2_RPT / 2; ++k) {
            const float ve = fmax3(pr[k], pr[k + 1], hm[2 * k + 4][c]);          // rows 2k   .. 2k+4
            const float vo = fmax3(hm[2 * k + 1][c], pr[k + 1], pr[k + 2]);      // rows 2k+1 .. 2k+5
            const float ce = c == 0 ? ctr[2 * k].x : c == 1 ? ctr[2 * k].y : c == 2 ? ctr[2 * k].z : ctr[2 * k].w;
            const float co = c == 0 ? ctr[2 * k + 1].x : c == 1 ? ctr[2 * k + 1].y : c == 2 ? ctr[2 * k + 1].z : ctr[2 * k + 1].w;
            if (ce == ve) peakmask |= 1u << ((2 * k) * 4 + c);
            if (co == vo) peakmask |= 1u << ((2 * k + 1) * 4 + c);
        }
    }

    // ---- 3x3 dilation of the one-hot bin masks, four pixels per word -------------------------------------------------
    // tbin row i <-> image row y0-1+i, byte k <-> image column x0-16+k; this thread needs columns xq-1 .. xq+4 = bytes 4*lane+15 .. +20
    unsigned hor[K2_RPT + 2];                   // per staged row: byte c = OR of one-hot(bin) over columns xq+c-1 .. xq+c+1
#pragma unroll
    for (int i = 0; i < K2_RPT + 2; ++i) {
        const uint8_t * row = &tbin[(ty0 + i) * K2_BW + lane * 4];
        const unsigned w1 = *reinterpret_cast<const unsigned *>(row + 16);
        const unsigned bl = row[15], br = row[20];
        // one-hot of the four bytes by a byte-table lookup: selector nibble c = bin of byte c
        const unsigned t = w1 | (w1 >> 4);
        const unsigned oh = __byte_perm(0x08040201u, 0x80402010u, __byte_perm(t, 0u, 0x4420u));
        // out-of-image staged bytes are 0 (bin 0); they are only ever combined into border pixels, which are forced to 0
        hor[i] = (oh | (oh << 8) | (oh >> 8)) | (1u << bl) | (0x01000000u << br);
    }
    unsigned colkeep = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) { const int x = xq + c; if (x >= 1 && x < P.w - 1) colkeep |= 0xFFu << (8 * c); }
    unsigned short * pmv = P.mv + slot * P.stride;
#pragma unroll
    for (int j = 0; j < K2_RPT; ++j) {
        const int y = y0 + ty0 + j;
        if (y < P.h && xq < P.wq) {
            const unsigned mword = hor[j] | hor[j + 1] | hor[j + 2];
            const bool vy = (y >= 1) && (y < P.h - 1);
            const unsigned mk = vy ? (mword & colkeep) : 0u;                  // four mask bytes -> four 16-bit entries, visited byte 0
            *reinterpret_cast<uint2 *>(pmv + (size_t)y * P.pitch + xq) = make_uint2(__byte_perm(mk, 0u, 0x4140u), __byte_perm(mk, 0u, 0x4342u));
        }
    }

    return peakmask;
}

// Append this thread's peaks to the frame's (unordered) seed list: one global atomic per CTA (K3 ranks exactly, so the
// list order is irrelevant).  *s_cnt must be 0 on entry; contains CTA barriers, so every thread of the CTA calls it.
template <int K2_RPT>
__device__ __forceinline__ void nms_append(unsigned peakmask, const float * tmag, int * s_cnt, int * s_base, const Planes & P, const Seeds & S,
                                           FrameCtrl * ctrl, int slot, int x0, int y0, float gmax, float thr, int lane, int warp)
{
    const int ty0 = warp * K2_RPT, xq = x0 + lane * 4;
    const int cnt = __popc(peakmask);
    int o = 0;
    if (cnt) o = atomicAdd(s_cnt, cnt);
    __syncthreads();
    const int total = *s_cnt;
    if (total == 0) return;                     // CTA-uniform
    if (threadIdx.x == 0) *s_base = atomicAdd(&ctrl[slot].n_seeds, total);
    __syncthreads();
    if (!cnt) return;
    o += *s_base;
    const float scale = __fdiv_rn((float)kSeedBuckets, __fsub_rn(gmax, thr));
    unsigned * us_mag = S.us_mag + (size_t)slot * S.cap; unsigned * us_idx = S.us_idx + (size_t)slot * S.cap;
    int * hist = S.hist + (size_t)slot * kSeedBuckets;
    int * rowhist = S.rowhist + (size_t)slot * S.rows;
    while (peakmask) {
        const int b = __ffs(peakmask) - 1;
        peakmask &= peakmask - 1;
        const int j = b >> 2, c = b & 3;
        if (o < S.cap) {
            const float m = tmag[(ty0 + 2 + j) * K2_MW + lane * 4 + 4 + c];
            us_mag[o] = __float_as_uint(m);
            us_idx[o] = (unsigned)((y0 + ty0 + j) * P.w + (xq + c));      // raster index in the dense W x H image
            atomicAdd(&hist[seed_bucket_scaled(m, thr, scale)], 1);
            atomicAdd(&rowhist[y0 + ty0 + j], 1);                         // peaks per image row: raster position of every peak (k_seed_rank)
        } else {
            ctrl[slot].seed_overflow = 1;
        }
        ++o;
    }
}

// one tile per CTA
template <int K2_RPT, int MINB>
__global__ void __launch_bounds__(32 * (K2_TH / K2_RPT), MINB)
k_nms5x5_seeds(const __grid_constant__ CUtensorMap tmap_mag, const __grid_constant__ CUtensorMap tmap_bin,
               Planes P, Seeds S, FrameCtrl * ctrl, int tma_x, int tma_y)
{
    __shared__ __align__(128) float tmag[K2_MH * K2_MW];
    __shared__ __align__(128) uint8_t tbin[K2_BH * K2_BW];
    __shared__ __align__(8) uint64_t bar;
    __shared__ int s_cnt, s_base;

    const int slot = blockIdx.z;
    const int x0 = blockIdx.x * K2_TW, y0 = blockIdx.y * K2_TH;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); s_cnt = 0; }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, K2_MH * K2_MW * 4 + K2_BH * K2_BW);
        tma_load_3d(tmag, &tmap_mag, x0 - 4 + tma_x, y0 - 2 + tma_y, slot, &bar);
        tma_load_3d(tbin, &tmap_bin, x0 - 16 + tma_x, y0 - 1 + tma_y, slot, &bar);
    }
    const float gmax = __uint_as_float(ctrl[slot].gmax_bits);
    const float thr = __fmul_rn(gmax, kSeedFactor);
    // mag >= 0 and thr >= 0: for a finite float m, m > thr <=> m >= tup, tup = the float after thr
    const float tup = __uint_as_float(__float_as_uint(thr) + 1u);
    mbar_wait(&bar, 0);
    const unsigned peakmask = nms_tile<K2_RPT>(tmag, tbin, P, slot, x0, y0, tup, lane, warp);
    nms_append<K2_RPT>(peakmask, tmag, &s_cnt, &s_base, P, S, ctrl, slot, x0, y0, gmax, thr, lane, warp);
}

// Persistent form (the default, LRB_K2_CFG=15: three CTAs per SM under a 150 KB shared-memory split): a resident grid walks the tiles of all frames; while a tile is being evaluated the copy engine already
// fetches the CTA's next one into the other stage (two mbarriers, phase = use count of the stage), so the HBM latency of
// a tile hides behind the arithmetic of the previous one instead of idling the CTA (the one-shot form spends half of its
// warp-cycles in the mbarrier wait: profiles/r1_k2_v2.md).  Measured on B200 (profiles/r1_k2_forms.txt): 111 vs 124 us
// per 32 frames (0.55 vs 0.49 of the HBM peak).  It must NOT be launched with the maximum shared-memory split (4 CTAs x 50 KB per
// SM): the SMs then keep a 28 KB L1 and the other engines' flood kernels, which share them, lose a fifth of their speed
// (whole job 3400 -> 2700 images/s) - hence the explicit 66 % preferred carveout in launch_nms.
constexpr int K2_BIN_BYTES = (K2_BH * K2_BW + 127) / 128 * 128;
constexpr int K2_STAGE_OFF_BIN = 2 * K2_MH * K2_MW * 4;
constexpr int K2_PERSIST_SMEM = K2_STAGE_OFF_BIN + 2 * K2_BIN_BYTES + 64 + 128;      // + barriers / counters + alignment slack
static_assert((K2_MH * K2_MW * 4) % 128 == 0, "TMA destinations must stay 128-byte aligned");

template <int K2_RPT, int MINB>
__global__ void __launch_bounds__(32 * (K2_TH / K2_RPT), MINB)
k_nms5x5_seeds_persistent(const __grid_constant__ CUtensorMap tmap_mag, const __grid_constant__ CUtensorMap tmap_bin,
                          Planes P, Seeds S, FrameCtrl * ctrl, int tma_x, int tma_y, int tiles_x, int tiles_y, int n_tiles)
{
    extern __shared__ unsigned char k2_smem_raw[];
    unsigned char * sm = k2_smem_raw + ((128u - (smem_u32(k2_smem_raw) & 127u)) & 127u);
    auto tmag_of = [&](int stage) { return reinterpret_cast<float *>(sm + stage * (K2_MH * K2_MW * 4)); };
    auto tbin_of = [&](int stage) { return sm + K2_STAGE_OFF_BIN + stage * K2_BIN_BYTES; };
    uint64_t * bar = reinterpret_cast<uint64_t *>(sm + K2_STAGE_OFF_BIN + 2 * K2_BIN_BYTES);          // bar[0], bar[1]
    int * s_cnt = reinterpret_cast<int *>(bar + 2);                                                    // s_cnt[0], s_cnt[1]
    int * s_base = s_cnt + 2;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int per_frame = tiles_x * tiles_y;
    auto issue = [&](int t, int stage) {        // thread 0 only
        const int slot = t / per_frame, r = t - slot * per_frame, by = r / tiles_x, bx = r - by * tiles_x;
        mbar_arrive_expect_tx(&bar[stage], K2_MH * K2_MW * 4 + K2_BH * K2_BW);
        tma_load_3d(tmag_of(stage), &tmap_mag, bx * K2_TW - 4 + tma_x, by * K2_TH - 2 + tma_y, slot, &bar[stage]);
        tma_load_3d(tbin_of(stage), &tmap_bin, bx * K2_TW - 16 + tma_x, by * K2_TH - 1 + tma_y, slot, &bar[stage]);
    };
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_fence_init();
        s_cnt[0] = 0; s_cnt[1] = 0;
        if ((int)blockIdx.x < n_tiles) issue(blockIdx.x, 0);
    }
    __syncthreads();
    int it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int stage = it & 1;
        // stage^1 was last read in iteration it-1, which ended with a CTA barrier
        if (threadIdx.x == 0) { if (t + (int)gridDim.x < n_tiles) issue(t + gridDim.x, stage ^ 1); s_cnt[stage ^ 1] = 0; }
        const int slot = t / per_frame, r = t - slot * per_frame, by = r / tiles_x, bx = r - by * tiles_x;
        const int x0 = bx * K2_TW, y0 = by * K2_TH;
        const float gmax = __uint_as_float(__ldg(&ctrl[slot].gmax_bits));
        const float thr = __fmul_rn(gmax, kSeedFactor);
        const float tup = __uint_as_float(__float_as_uint(thr) + 1u);
        mbar_wait(&bar[stage], (it >> 1) & 1);
        const float * tm = tmag_of(stage);
        const unsigned peakmask = nms_tile<K2_RPT>(tm, tbin_of(stage), P, slot, x0, y0, tup, lane, warp);
        nms_append<K2_RPT>(peakmask, tm, &s_cnt[stage], s_base, P, S, ctrl, slot, x0, y0, gmax, thr, lane, warp);
        __syncthreads();                        // everyone is done with this stage (and with s_base)
    }
}

// exclusive scan of the bucket histogram; one block of 1024 threads per frame, 4 buckets per thread
__global__ void __launch_bounds__(1024) k_seed_scan(Seeds S, FrameCtrl * ctrl)
{
    static_assert(kSeedBuckets == 4096, "4 buckets per thread");
    __shared__ int wsum[32];
    const int slot = blockIdx.x;
    int * hist = S.hist + (size_t)slot * kSeedBuckets; int * start = S.start + (size_t)slot * kSeedBuckets;
    int * cursor = S.cursor + (size_t)slot * kSeedBuckets;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int4 h = *reinterpret_cast<const int4 *>(hist + 4 * t);
    const int mine = h.x + h.y + h.z + h.w;
    int incl = mine;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += v; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int v = wsum[lane];
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, s); if (lane >= s) v += u; }
        wsum[lane] = v;
    }
    __syncthreads();
    const int ex = incl - mine + (warp ? wsum[warp - 1] : 0);
    *reinterpret_cast<int4 *>(start + 4 * t) = make_int4(ex, ex + h.x, ex + h.x + h.y, ex + h.x + h.y + h.z);
    *reinterpret_cast<int4 *>(cursor + 4 * t) = make_int4(0, 0, 0, 0);
    {   // exclusive scan of the per-row peak counts: rowstart[r] = raster position of the first peak of row r
        __syncthreads();                        // wsum is reused
        const int rows = S.rows, per = (rows + 1023) / 1024;
        const int * rh = S.rowhist + (size_t)slot * rows; int * rs = S.rowstart + (size_t)slot * rows; int * rc = S.rowcursor + (size_t)slot * rows;
        const int r0 = t * per, r1 = min(rows, r0 + per);
        int sum = 0;
        for (int r = r0; r < r1; ++r) sum += rh[r];
        int inc = sum;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += v; }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int v = wsum[lane];
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, s); if (lane >= s) v += u; }
            wsum[lane] = v;
        }
        __syncthreads();
        int run = inc - sum + (warp ? wsum[warp - 1] : 0);
        for (int r = r0; r < r1; ++r) { rs[r] = run; run += rh[r]; rc[r] = 0; }
    }
    if (t == 0) {
        if (ctrl[slot].n_seeds > S.cap) { ctrl[slot].n_seeds = S.cap; ctrl[slot].seed_overflow = 1; }
        const int ns = ctrl[slot].n_seeds;
        ctrl[slot].live[0] = ns;
        ctrl[slot].win_lo = 0; ctrl[slot].win_hi = min(ns, min(max(S.win_min, ns / S.win_div), S.win_first));
        ctrl[slot].win_live = 0; ctrl[slot].win_first = 0x7fffffff; ctrl[slot].win_blocks = 0;
    }
}

__global__ void __launch_bounds__(256) k_seed_scatter(Seeds S, FrameCtrl * ctrl)
{
    const int slot = blockIdx.y;
    const int n = min(ctrl[slot].n_seeds, S.cap);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gmax = __uint_as_float(ctrl[slot].gmax_bits);
    const float thr = __fmul_rn(gmax, kSeedFactor);
    const size_t so = (size_t)slot * S.cap, bo = (size_t)slot * kSeedBuckets;
    const unsigned m = S.us_mag[so + i];
    const int d = seed_bucket(__uint_as_float(m), thr, gmax);
    const int pos = S.start[bo + d] + atomicAdd(&S.cursor[bo + d], 1);
    const unsigned ix = S.us_idx[so + i];
    S.b_mag[so + pos] = m;
    S.b_idx[so + pos] = ix;
    const int row = (int)(ix / (unsigned)S.roww), col = (int)(ix - (unsigned)row * (unsigned)S.roww);
    const size_t ro = (size_t)slot * S.rows + row;
    S.rb_col[so + S.rowstart[ro] + atomicAdd(&S.rowcursor[ro], 1)] = col;
}

// exact rank inside the bucket by counting, then per-seed flood parameters
__global__ void __launch_bounds__(256) k_seed_rank(Planes P, Seeds S, FrameCtrl * ctrl)
{
    const int slot = blockIdx.y;
    const int n = min(ctrl[slot].n_seeds, S.cap);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gmax = __uint_as_float(ctrl[slot].gmax_bits);
    const float thr = __fmul_rn(gmax, kSeedFactor);
    const size_t so = (size_t)slot * S.cap, bo = (size_t)slot * kSeedBuckets;
    const unsigned m = S.b_mag[so + i], ix = S.b_idx[so + i];
    const int d = seed_bucket(__uint_as_float(m), thr, gmax);
    const int lo = S.start[bo + d], hi = lo + S.hist[bo + d];
    // the in-bucket rank costs O(bucket^2); an input with tens of thousands of (nearly) equal magnitudes is rejected
    // loudly instead of occupying the GPU for seconds
    if (hi - lo > 16384) { ctrl[slot].seed_overflow = 2; return; }
    int greater = 0, eq_before = 0, eq_total = 0;
    for (int j = lo; j < hi; ++j) {
        const unsigned mj = S.b_mag[so + j], ij = S.b_idx[so + j];
        greater += (mj > m);                                  // positive floats: bit order == value order
        eq_total += (mj == m); eq_before += (mj == m && ij < ix);
    }
    const int key = lo + greater;                             // dense rank of the magnitude: exact ties share it
    const int rank = key + eq_before;                         // (magnitude desc, raster asc); final unless the peak is tied (k_seed_replay)
    const int r = ix / P.w, c = ix - r * P.w;
    {   // the peak's place in the reference's peak list before the sort = its raster position (filter.cpp:174-186, single thread)
        const size_t ro = (size_t)slot * S.rows + r;
        const int rs = S.rowstart[ro], re = rs + S.rowhist[ro];
        int rp = rs;
        for (int j = rs; j < re; ++j) rp += (S.rb_col[so + j] < c);
        const unsigned tied = eq_total > 1 ? 0x80000000u : 0u;
        S.rp_el[so + rp] = make_uint2((unsigned)key, tied | (unsigned)rank);
        if (tied) atomicAdd(&ctrl[slot].n_tied, 1);
    }
    const size_t po = (size_t)slot * P.stride + (size_t)r * P.pitch + c;
    const int b = P.bin[po];
    S.idx[so + rank] = ix;
    S.bin[so + rank] = (uint8_t)b;
    // flood threshold (1 - 0.25f) * G_b(seed)   (reference src/filter.cpp:119-121)
    const float2 gseed = P.dxy[po];
    S.thr[so + rank] = __fmul_rn(kTraceKeep, orient_energy(gseed.x, gseed.y, b));
    S.status[so + rank] = kStatusLive;
    S.fail_epoch[so + rank] = 0u;
    S.size[so + rank] = 0;
}

// zero the per-frame control block and the bucket histogram
__global__ void __launch_bounds__(256) k_frame_init(Seeds S, FrameCtrl * ctrl)
{
    const int slot = blockIdx.x;
    int * c = reinterpret_cast<int *>(&ctrl[slot]);
    for (int i = threadIdx.x; i < (int)(sizeof(FrameCtrl) / sizeof(int)); i += blockDim.x) c[i] = 0;
    int * hist = S.hist + (size_t)slot * kSeedBuckets;
    for (int i = threadIdx.x; i < kSeedBuckets; i += blockDim.x) hist[i] = 0;
    int * rowhist = S.rowhist + (size_t)slot * S.rows;
    for (int i = threadIdx.x; i < S.rows; i += blockDim.x) rowhist[i] = 0;
}

} // namespace

void launch_frame_init(const Seeds & S, FrameCtrl * ctrl, int n_slots, cudaStream_t st)
{
    k_frame_init<<<n_slots, 256, 0, st>>>(S, ctrl);
}

void launch_nms(const CUtensorMap & tmap_mag, const CUtensorMap & tmap_bin, const Planes & P, const Seeds & S,
                FrameCtrl * ctrl, int n_slots, int tma_x, int tma_y, cudaStream_t st)
{
    dim3 grid((P.w + K2_TW - 1) / K2_TW, (P.h + K2_TH - 1) / K2_TH, n_slots);
    static const int cfg = [] { const char * v = getenv("LRB_K2_CFG"); return v ? atoi(v) : 15; }();
    if (cfg >= 10) {                            // persistent, double-buffered (default: three CTAs per SM under a 66 % split)
        auto kern = k_nms5x5_seeds_persistent<8, 3>;
        const int threads = 128;
        static PerDeviceOnce once; static int per_sm[PerDeviceOnce::kMaxDev], n_sm[PerDeviceOnce::kMaxDev]; static cudaError_t attr_err[PerDeviceOnce::kMaxDev];
        const int d = once.run([&] {
            const int dv = PerDeviceOnce::dev();
            cudaError_t e1 = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K2_PERSIST_SMEM);
            // a kernel that asks for the maximum split leaves the SMs with a 28 KB L1, and the other engines' flood kernels,
            // which share those SMs, lose a fifth of their speed: three CTAs per SM under a 66 % split
            cudaError_t e2 = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 66);
            attr_err[dv] = e1 != cudaSuccess ? e1 : e2;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm[dv], kern, threads, K2_PERSIST_SMEM) != cudaSuccess || per_sm[dv] < 1) per_sm[dv] = 1;
            n_sm[dv] = 148; cudaDeviceGetAttribute(&n_sm[dv], cudaDevAttrMultiProcessorCount, dv);
        });
        if (attr_err[d] != cudaSuccess) fprintf(stderr, "[lrb] K2: shared-memory opt-in failed on device %d: %s (the launch below will fail)\n", d, cudaGetErrorString(attr_err[d]));
        const int n_tiles = (int)(grid.x * grid.y * grid.z);
        const int use_per_sm = per_sm[d] < 3 ? per_sm[d] : 3;
        const int blocks = n_tiles < n_sm[d] * use_per_sm ? n_tiles : n_sm[d] * use_per_sm;
        LRB_CARVEOUT(kern);
        kern<<<blocks, threads, K2_PERSIST_SMEM, st>>>(tmap_mag, tmap_bin, P, S, ctrl, tma_x, tma_y, (int)grid.x, (int)grid.y, n_tiles);
        return;
    }
    // LRB_K2_CFG=2: the one-shot form (one tile per CTA) — the documented alternative, see profiles/r1_sweeps.md
    LRB_CARVEOUT((k_nms5x5_seeds<8, 6>));
    k_nms5x5_seeds<8, 6><<<grid, 128, 0, st>>>(tmap_mag, tmap_bin, P, S, ctrl, tma_x, tma_y);
}

void launch_seed_order(const Planes & P, const Seeds & S, FrameCtrl * ctrl, int n_slots, int seeds_hint, cudaStream_t st)
{
    LRB_CARVEOUT(k_seed_scan); LRB_CARVEOUT(k_seed_scatter); LRB_CARVEOUT(k_seed_rank);
    k_seed_scan<<<n_slots, 1024, 0, st>>>(S, ctrl);
    dim3 grid((seeds_hint + 255) / 256, n_slots);
    k_seed_scatter<<<grid, 256, 0, st>>>(S, ctrl);
    k_seed_rank<<<grid, 256, 0, st>>>(P, S, ctrl);
    launch_seed_replay(P, S, ctrl, n_slots, -1, true, st);
}

void nms_tile_dims(int * mw, int * mh, int * bw, int * bh) { *mw = K2_MW; *mh = K2_MH; *bw = K2_BW; *bh = K2_BH; }

} // namespace lrb
