// k6_ransac.cu — K6..K9: segment filter + pencil model, the fixed-seed Knuth sampler, the hypotheses x segments
// scoring grid, and refit + inlier/garbage assignment for up to four vanishing-point structures.
//
// Replaces _filter_lines (reference src/interface.cpp:26-32), bounding_box / normalize_lines (src/geometry.cpp:96-112,
// 258-282), the LinePencilModel constructor (src/line_pencil.cpp:25-32), choice_knuth (src/math_utils.cpp:14-39),
// RANSAC_Estimator::solve (src/estimator.h:37-78), sample_check / fit / inlier_score / error / fit_optimal
// (src/line_pencil.cpp:89-140), inclination (src/geometry.cpp:214-229) and estimate_multiple_structures
// (src/estimator.h:99-145).
//
// Everything that decides group membership is evaluated in the reference's float32 operation order: per-hypothesis
// scores are sequential sums over the observation list, the winner is the first strict maximum in draw order, and the
// refit scatter matrix is accumulated sequentially.  Tensor cores do not apply (inner dimension 3); the grid runs on the
// FP32 pipe, one thread per hypothesis with the observations broadcast from shared memory.
//
// The sampler reproduces std::mt19937 + libstdc++ generate_canonical<float,24> + Algorithm S exactly.  The uniform
// stream depends only on the seed, so it is generated once per seed (mt_stream.cu) as k = 2^32 * u (an exact integer,
// multiple of 256) and shared by every frame; "select" <=> (n - t) * k < (2 - m) * 2^32 in integer arithmetic.
#include "common.cuh"
#include "eig.cuh"
#include "kernels.h"
#include "pencil.cuh"

namespace lrb {

namespace {

// ---- K6: filter + bounding box + pencil model --------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_filter_model(Comps K, Lines L, FrameCtrl * ctrl, float min_length, int rounds_total, const float * refined, size_t refined_stride)
{
    __shared__ int wsum[32];
    __shared__ float red[4][32];
    __shared__ int s_base;
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const size_t ko = (size_t)slot * K.cap, lo = (size_t)slot * L.cap;
    // with `refine` the filter reads the merged segments (k_refine.cu); the "fewer than two raw segments" exit
    // (interface.cpp:50-54) is taken on the unmerged count either way
    const int ncomp = refined ? C.n_merged : C.n_comp;
    SegRec * seg = reinterpret_cast<SegRec *>(L.seg) + lo;
    const SegRec * raw = refined ? reinterpret_cast<const SegRec *>(refined + (size_t)slot * refined_stride)
                                 : reinterpret_cast<const SegRec *>(K.seg) + ko;
    const bool floods_complete = (C.live[rounds_total] == 0 || C.flood_done) && !C.flood_overflow;
    if (t == 0) {
        s_base = 0;
        // (re)start the grouping state: this kernel is also the entry point of a repair pass
        for (int k = 0; k < kMaxModels; ++k) { C.best[k] = 0ull; C.winner[k] = -1; C.n_obs[k] = 0; }
        C.stream_pos = 0ull; C.stream_overflow = 0; C.n_solves = 0; C.samp_done = 0;
        if (floods_complete) C.fit_done = 1;             // the fit kernels ran on final floods: never redo them (they consume the owner stamps)
    }
    __syncthreads();
    if (C.n_comp < 2 || !floods_complete || (refined && C.refine_overflow)) {                       // interface.cpp:50-54
        if (t == 0) { C.n_lines = 0; C.num_obs = 0; }
        return;
    }
    const float ml = fmaxf(min_length, kLineMinLength);                         // interface.cpp:29
    float mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY;
    for (int i0 = 0; i0 < ncomp; i0 += 1024) {
        const int i = i0 + t;
        int keep = 0; SegRec s;
        if (i < ncomp) {
            s = raw[i];
            const float x = __fsub_rn(s.x2, s.x1), y = __fsub_rn(s.y2, s.y1);
            const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)));      // geometry.cpp:90-93
            keep = (len > ml) && (s.err < kLineMaxErr);
        }
        int total;
        const int base = s_base;                 // read before the barriers inside the scan: thread 0 advances s_base right after it
        const int pos = base + block_scan_flag(keep, wsum, total);
        if (keep) {
            if (pos < L.cap) {
                seg[pos] = s;
                mnx = fminf(mnx, fminf(s.x1, s.x2)); mxx = fmaxf(mxx, fmaxf(s.x1, s.x2));
                mny = fminf(mny, fminf(s.y1, s.y2)); mxy = fmaxf(mxy, fmaxf(s.y1, s.y2));
            } else C.line_overflow = 1;
        }
        if (t == 0) s_base += total;
        __syncthreads();
    }
    const int N = min(s_base, L.cap);
    // bounding box of all endpoints (geometry.cpp:96-112): min/max are exact, any order
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, s)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, s));
        mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, s)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, s));
    }
    if (lane == 0) { red[0][warp] = mnx; red[1][warp] = mny; red[2][warp] = mxx; red[3][warp] = mxy; }
    __syncthreads();
    mnx = red[0][0]; mny = red[1][0]; mxx = red[2][0]; mxy = red[3][0];
    for (int w = 1; w < 32; ++w) { mnx = fminf(mnx, red[0][w]); mny = fminf(mny, red[1][w]); mxx = fmaxf(mxx, red[2][w]); mxy = fmaxf(mxy, red[3][w]); }
    const float sx = __fsub_rn(mxx, mnx), sy = __fsub_rn(mxy, mny);
    const float cx = __fadd_rn(mnx, __fmul_rn(0.5f, sx)), cy = __fadd_rn(mny, __fmul_rn(0.5f, sy));   // geometry.cpp:272-276
    const float scale = fmaxf(sx, sy);                                                                // bbox_size().maxCoeff()
    if (t == 0) {
        C.n_lines = N; C.num_obs = N; C.n_obs[0] = N; C.n_solves = 0;
        C.bbox[0] = mnx; C.bbox[1] = mny; C.bbox[2] = mxx; C.bbox[3] = mxy; C.cx = cx; C.cy = cy; C.scale = scale;
    }
    for (int i = t; i < N; i += 1024) {
        const SegRec s = seg[i];
        // normalize_lines geometry.cpp:258-270
        const float x1 = __fdiv_rn(__fsub_rn(s.x1, cx), scale), y1 = __fdiv_rn(__fsub_rn(s.y1, cy), scale);
        const float x2 = __fdiv_rn(__fsub_rn(s.x2, cx), scale), y2 = __fdiv_rn(__fsub_rn(s.y2, cy), scale);
        // homogeneous geometry.cpp:64-69: (x1,y1,1) x (x2,y2,1), then unit length
        const float hx = __fsub_rn(__fmul_rn(y1, 1.f), __fmul_rn(1.f, y2));
        const float hy = __fsub_rn(__fmul_rn(1.f, x2), __fmul_rn(x1, 1.f));
        const float hz = __fsub_rn(__fmul_rn(x1, y2), __fmul_rn(y1, x2));
        const float nrm = __fsqrt_rn(sum3(__fmul_rn(hx, hx), __fmul_rn(hy, hy), __fmul_rn(hz, hz)));
        L.h[(lo + i) * 3 + 0] = __fdiv_rn(hx, nrm); L.h[(lo + i) * 3 + 1] = __fdiv_rn(hy, nrm); L.h[(lo + i) * 3 + 2] = __fdiv_rn(hz, nrm);
        L.anchor[(lo + i) * 2 + 0] = __fdiv_rn(__fadd_rn(x2, x1), 2.f); L.anchor[(lo + i) * 2 + 1] = __fdiv_rn(__fadd_rn(y2, y1), 2.f);
        const float ddx = __fsub_rn(x2, x1), ddy = __fsub_rn(y2, y1);
        const float ln = __fsqrt_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)));
        L.len[lo + i] = ln;
        L.dir[(lo + i) * 2 + 0] = __fdiv_rn(ddx, ln); L.dir[(lo + i) * 2 + 1] = __fdiv_rn(ddy, ln);
        L.inlier[lo + i] = -1; L.garbage[lo + i] = 0; L.obs[lo + i] = i;
    }
}

// decision thresholds of the fast inlier test in k_score: (1 - tol +- 4e-6)^2, evaluated in double
constexpr double kInTheta = 1.0 - (double)kTolInlier;
constexpr float kInHi = (float)((kInTheta + 4e-6) * (kInTheta + 4e-6)), kInLo = (float)((kInTheta - 4e-6) * (kInTheta - 4e-6));

// ---- K8: scoring grid -------------------------------------------------------------------------------------------------
constexpr int SC_THREADS = 128;
constexpr int SC_CHUNK = 1024;

__global__ void __launch_bounds__(SC_THREADS) k_score(Lines L, FrameCtrl * ctrl, int solve, int T)
{
    __shared__ float4 s_ad[SC_CHUNK];      // anchor.x, anchor.y, dir.x, dir.y
    __shared__ float s_len[SC_CHUNK];
    __shared__ unsigned long long s_best[SC_THREADS / 32];
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    const int n = C.n_obs[solve];
    const size_t lo = (size_t)slot * L.cap;
    const int * obs = L.obs + lo;
    const int i = blockIdx.x * SC_THREADS + threadIdx.x;           // draw index
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    bool ok = false; Hyp h = make_hyp(0.f, 0.f, 0.f);
    if (i < T) {
        const int * pr = L.pairs + (size_t)slot * 2 * T + 2 * i;
        const int a = obs[pr[0]], b = obs[pr[1]];
        const float * A = L.h + (lo + a) * 3; const float * B = L.h + (lo + b) * 3;
        const float ex = __fsub_rn(A[0], B[0]), ey = __fsub_rn(A[1], B[1]), ez = __fsub_rn(A[2], B[2]);
        ok = __fsqrt_rn(sum3(__fmul_rn(ex, ex), __fmul_rn(ey, ey), __fmul_rn(ez, ez))) > kDegeneracyTol;   // line_pencil.cpp:89-98
        float p0, p1, p2; cross3(A, B, p0, p1, p2);                                                     // line_pencil.cpp:101-108
        h = make_hyp(p0, p1, p2);
    }
    float score = 0.f;
    for (int c0 = 0; c0 < n; c0 += SC_CHUNK) {
        const int cn = min(SC_CHUNK, n - c0);
        __syncthreads();
        for (int k = threadIdx.x; k < cn; k += SC_THREADS) {
            const int q = obs[c0 + k];
            const float2 an = *reinterpret_cast<const float2 *>(L.anchor + (lo + q) * 2);
            const float2 di = *reinterpret_cast<const float2 *>(L.dir + (lo + q) * 2);
            s_ad[k] = make_float4(an.x, an.y, di.x, di.y);
            s_len[k] = L.len[lo + q];
        }
        __syncthreads();
        if (ok) {
#pragma unroll 4
            for (int k = 0; k < cn; ++k) {
                const float4 v = s_ad[k];
                // inlier <=> 1 - |v^ . d| < tol (line_pencil.cpp:137-140).  Only this boolean enters the score, and for
                // all but a ~1e-5 sliver of pairs it is decided without the IEEE sqrt and the two divisions:
                // (v.d)^2 against |v|^2 (1 - tol -+ delta)^2, delta = 4e-6 — five times the worst-case sum of the
                // rounding errors of both evaluations — the sliver takes the exact path.
                const float vx = h.ideal ? h.px : __fadd_rn(-v.x, h.px), vy = h.ideal ? h.py : __fadd_rn(-v.y, h.py);
                const float D = __fmaf_rn(vx, v.z, __fmul_rn(vy, v.w));       // the fast test may contract: fewer roundings, inside the same delta
                const float Z = __fmaf_rn(vx, vx, __fmul_rn(vy, vy));
                const float D2 = __fmul_rn(D, D);
                bool in = D2 > __fmul_rn(Z, kInHi);
                if (!in && !(D2 < __fmul_rn(Z, kInLo))) in = line_error(h, v.x, v.y, v.z, v.w) < kTolInlier;
                if (in) score = __fadd_rn(score, s_len[k]);                  // adding the +0 of an outlier changes nothing
            }
        }
    }
    if (i < T) L.scores[((size_t)slot * kMaxModels + solve) * T + i] = ok ? score : -1.f;
    // first strict maximum in draw order (estimator.h:62-70): max score, then lowest draw index
    unsigned long long key = (ok && score > 0.f) ? (((unsigned long long)__float_as_uint(score) << 32) | (unsigned)(0xFFFFFFFFu - (unsigned)i)) : 0ull;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, key, s); key = o > key ? o : key; }
    if (lane == 0) s_best[warp] = key;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < SC_THREADS / 32; ++w) key = s_best[w] > key ? s_best[w] : key;
        if (key) atomicMax(&C.best[solve], key);
    }
}

// ---- K8, latency form (one or two frames per launch) ----------------------------------------------------------------------
// A CTA scores 32 hypotheses (lane = hypothesis) with SCW_WARPS warps.  Only the BOOLEAN "inlier" enters a score, and it is
// order-free; the sum of the inliers' lengths is a float sum in observation order (estimator.h:52-60) and is not.  So the warps
// share a staged chunk of observations: each tests its slice for the 32 hypotheses and leaves one bit per (hypothesis,
// observation) in shared memory, and warp 0 then walks the bits of its hypothesis in observation order and adds the lengths —
// the same sequential sum, with SCW_WARPS times the parallelism on the part that costs (one warp per 32 hypotheses meant 313
// warps on the whole GPU for a frame: 7.4 ms for the 35 000 observations of a 4K clutter frame, 0.2 ms of a 1080p frame's latency).
constexpr int SCW_WARPS = 8;
constexpr int SCW_THREADS = 32 * SCW_WARPS;
constexpr int SCW_SEG = SC_CHUNK / SCW_WARPS;            // observations of a chunk tested by one warp
static_assert(SCW_SEG % 32 == 0, "a warp's slice is whole mask words");

__global__ void __launch_bounds__(SCW_THREADS) k_score_wide(Lines L, FrameCtrl * ctrl, int solve, int T)
{
    __shared__ float4 s_ad[SC_CHUNK];      // anchor.x, anchor.y, dir.x, dir.y
    __shared__ float s_len[SC_CHUNK];
    __shared__ unsigned s_mask[SC_CHUNK / 32][32];     // [word of the chunk][hypothesis]: inlier bits
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    const int n = C.n_obs[solve];
    const size_t lo = (size_t)slot * L.cap;
    const int * obs = L.obs + lo;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;                          // draw index (the same 32 hypotheses in every warp)

    bool ok = false; Hyp h = make_hyp(0.f, 0.f, 0.f);
    if (i < T) {
        const int * pr = L.pairs + (size_t)slot * 2 * T + 2 * i;
        const int a = obs[pr[0]], b = obs[pr[1]];
        const float * A = L.h + (lo + a) * 3; const float * B = L.h + (lo + b) * 3;
        const float ex = __fsub_rn(A[0], B[0]), ey = __fsub_rn(A[1], B[1]), ez = __fsub_rn(A[2], B[2]);
        ok = __fsqrt_rn(sum3(__fmul_rn(ex, ex), __fmul_rn(ey, ey), __fmul_rn(ez, ez))) > kDegeneracyTol;   // line_pencil.cpp:89-98
        float p0, p1, p2; cross3(A, B, p0, p1, p2);                                                     // line_pencil.cpp:101-108
        h = make_hyp(p0, p1, p2);
    }
    float score = 0.f;
    for (int c0 = 0; c0 < n; c0 += SC_CHUNK) {
        const int cn = min(SC_CHUNK, n - c0);
        __syncthreads();
        for (int k = threadIdx.x; k < cn; k += SCW_THREADS) {
            const int q = obs[c0 + k];
            const float2 an = *reinterpret_cast<const float2 *>(L.anchor + (lo + q) * 2);
            const float2 di = *reinterpret_cast<const float2 *>(L.dir + (lo + q) * 2);
            s_ad[k] = make_float4(an.x, an.y, di.x, di.y);
            s_len[k] = L.len[lo + q];
        }
        __syncthreads();
        // this warp's slice of the chunk: one mask word per 32 observations
#pragma unroll 1
        for (int w = 0; w < SCW_SEG / 32; ++w) {
            const int k0 = warp * SCW_SEG + w * 32;
            unsigned m = 0u;
            if (ok && k0 < cn) {
                const int kn = min(32, cn - k0);
#pragma unroll 4
                for (int j = 0; j < kn; ++j) {
                    const float4 v = s_ad[k0 + j];
                    // inlier <=> 1 - |v^ . d| < tol (line_pencil.cpp:137-140).  For all but a ~1e-5 sliver of pairs it is decided
                    // without the IEEE sqrt and the two divisions: (v.d)^2 against |v|^2 (1 - tol -+ delta)^2, delta = 4e-6 — five
                    // times the worst-case sum of the rounding errors of both evaluations — the sliver takes the
                    // exact path.
                    const float vx = h.ideal ? h.px : __fadd_rn(-v.x, h.px), vy = h.ideal ? h.py : __fadd_rn(-v.y, h.py);
                    const float D = __fmaf_rn(vx, v.z, __fmul_rn(vy, v.w));       // the fast test may contract: fewer roundings, inside the same delta
                    const float Z = __fmaf_rn(vx, vx, __fmul_rn(vy, vy));
                    const float D2 = __fmul_rn(D, D);
                    bool in = D2 > __fmul_rn(Z, kInHi);
                    if (!in && !(D2 < __fmul_rn(Z, kInLo))) in = line_error(h, v.x, v.y, v.z, v.w) < kTolInlier;
                    m |= in ? (1u << j) : 0u;
                }
            }
            s_mask[warp * (SCW_SEG / 32) + w][lane] = m;
        }
        __syncthreads();
        // the sequential sum, in observation order (adding the +0 of an outlier would change nothing, so outliers are skipped)
        if (warp == 0 && ok) {
            const int nw = (cn + 31) / 32;
            for (int w = 0; w < nw; ++w) {
                unsigned m = s_mask[w][lane];
                while (m) { const int j = __ffs(m) - 1; m &= m - 1; score = __fadd_rn(score, s_len[w * 32 + j]); }
            }
        }
    }
    if (warp != 0) return;
    if (i < T) L.scores[((size_t)slot * kMaxModels + solve) * T + i] = ok ? score : -1.f;
    // first strict maximum in draw order (estimator.h:62-70): max score, then lowest draw index
    unsigned long long key = (ok && score > 0.f) ? (((unsigned long long)__float_as_uint(score) << 32) | (unsigned)(0xFFFFFFFFu - (unsigned)i)) : 0ull;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, key, s); key = o > key ? o : key; }
    if (lane == 0 && key) atomicMax(&C.best[solve], key);
}

// ---- K9: refit on the winner's inliers, assign inliers / garbage, build the next observation list ----------------------
constexpr int kRefitChunk = 1024;       // terms of the six scatter sums staged per pass
__global__ void __launch_bounds__(1024) k_refit_assign(Lines L, FrameCtrl * ctrl, int solve, int T)
{
    __shared__ int wsum[32];
    __shared__ float s_cov[9];
    __shared__ float s_h[3];
    __shared__ int s_cnt[2];
    __shared__ float s_terms[6][kRefitChunk];
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    const int t = threadIdx.x;
    const int n = C.n_obs[solve];
    const int N = C.n_lines;
    const size_t lo = (size_t)slot * L.cap;
    int * obs = L.obs + lo; int * inl = L.tmp + lo;

    // winner -> best_h (recomputed from its pair: same arithmetic as in the scoring kernel)
    float b0 = 0.f, b1 = 0.f, b2 = 0.f; int win = -1;
    {
        const unsigned long long key = C.best[solve];
        if (key) {
            win = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
            const int * pr = L.pairs + (size_t)slot * 2 * T + 2 * win;
            cross3(L.h + (lo + obs[pr[0]]) * 3, L.h + (lo + obs[pr[1]]) * 3, b0, b1, b2);
        }
    }
    const Hyp hb = make_hyp(b0, b1, b2);
    // inliers of best_h in observation order (estimator.h:73-76)
    int ni = 0;
    for (int k0 = 0; k0 < n; k0 += 1024) {
        const int k = k0 + t; int f = 0, q = 0;
        if (k < n) {
            q = obs[k];
            f = line_error(hb, L.anchor[(lo + q) * 2], L.anchor[(lo + q) * 2 + 1], L.dir[(lo + q) * 2], L.dir[(lo + q) * 2 + 1]) < kTolInlier;
        }
        int total; const int pos = ni + block_scan_flag(f, wsum, total);
        if (f) inl[pos] = q;
        ni += total;
    }
    __syncthreads();
    // fit_optimal (line_pencil.cpp:111-128): empty index list => ALL lines of the model
    const bool all = (ni == 0);
    const int cnt = all ? N : ni;
    {
        // lower triangle (r >= c) of h^T diag(len) h, each entry a SEQUENTIAL sum of (h_r * len) * h_c in list order (that is the
        // reference's summation order).  The products are independent: all threads compute them into shared memory, chunk by chunk,
        // and six lanes only add — the running sums then wait on shared memory instead of on a chain of dependent global loads
        // (107 -> ~15 us per solve for a single frame).
        const int r = (t == 0) ? 0 : (t < 3 ? 1 : 2);
        const int c = (t == 0) ? 0 : (t == 1 ? 0 : (t == 2 ? 1 : t - 3));
        float acc = 0.f;
        for (int c0 = 0; c0 < cnt; c0 += kRefitChunk) {
            const int cn = min(kRefitChunk, cnt - c0);
            for (int k = t; k < cn; k += 1024) {
                const int q = all ? c0 + k : inl[c0 + k];
                const float h0 = L.h[(lo + q) * 3], h1 = L.h[(lo + q) * 3 + 1], h2 = L.h[(lo + q) * 3 + 2], ln = L.len[lo + q];
                const float a0 = __fmul_rn(h0, ln), a1 = __fmul_rn(h1, ln), a2 = __fmul_rn(h2, ln);
                s_terms[0][k] = __fmul_rn(a0, h0);
                s_terms[1][k] = __fmul_rn(a1, h0); s_terms[2][k] = __fmul_rn(a1, h1);
                s_terms[3][k] = __fmul_rn(a2, h0); s_terms[4][k] = __fmul_rn(a2, h1); s_terms[5][k] = __fmul_rn(a2, h2);
            }
            __syncthreads();
            if (t < 6) {
                const float * row = s_terms[t];
                int k = 0;
                if (c0 == 0) { acc = row[0]; k = 1; }
                for (; k < cn; ++k) acc = __fadd_rn(acc, row[k]);
            }
            __syncthreads();
        }
        if (t < 6) { s_cov[r * 3 + c] = acc; s_cov[c * 3 + r] = acc; }
    }
    __syncthreads();
    if (t == 0) {
        float eval[3], evec[9];
        sym_eig<3>(s_cov, eval, evec);
        int k = 0;
        for (int i = 1; i < 3; ++i) if (eval[i] < eval[k]) k = i;                 // minCoeff: first minimum
        s_h[0] = evec[k * 3]; s_h[1] = evec[k * 3 + 1]; s_h[2] = evec[k * 3 + 2];
        C.winner[solve] = win;
        C.best_h[solve][0] = b0; C.best_h[solve][1] = b1; C.best_h[solve][2] = b2;
        C.refined[solve][0] = s_h[0]; C.refined[solve][1] = s_h[1]; C.refined[solve][2] = s_h[2];
        s_cnt[0] = 0; s_cnt[1] = 0;
    }
    __syncthreads();
    const Hyp hr = make_hyp(s_h[0], s_h[1], s_h[2]);
    // estimator.h:127-138: inliers get the structure index, near misses become garbage
    int n_in = 0, n_gb = 0;
    for (int k = t; k < n; k += 1024) {
        const int q = obs[k];
        const float e = line_error(hr, L.anchor[(lo + q) * 2], L.anchor[(lo + q) * 2 + 1], L.dir[(lo + q) * 2], L.dir[(lo + q) * 2 + 1]);
        if (e < kTolInlier) { L.inlier[lo + q] = solve; ++n_in; }
        else if (e < kTolGarbage) { L.garbage[lo + q] = 1; ++n_gb; }
    }
    atomicAdd(&s_cnt[0], n_in); atomicAdd(&s_cnt[1], n_gb);
    __syncthreads();
    // next observation list: nonzero(inlier < 0 && garbage == 0) in index order (estimator.h:124, utils.h:16-30)
    int no = 0;
    for (int i0 = 0; i0 < N; i0 += 1024) {
        const int i = i0 + t; int f = 0;
        if (i < N) f = (L.inlier[lo + i] < 0) && (L.garbage[lo + i] == 0);
        int total; const int pos = no + block_scan_flag(f, wsum, total);
        if (f) obs[pos] = i;
        no += total;
    }
    if (t == 0) {
        C.num_obs -= s_cnt[0] + s_cnt[1];
        C.n_solves = solve + 1;
        if (solve + 1 < kMaxModels) C.n_obs[solve + 1] = no;
    }
}

// group ids as the reference reports them: garbage lines end up unassigned (estimator.h:142)
__global__ void k_finalize(Lines L, FrameCtrl * ctrl)
{
    const int slot = blockIdx.y;
    const int N = ctrl[slot].n_lines;
    const size_t lo = (size_t)slot * L.cap;
    SegRec * seg = reinterpret_cast<SegRec *>(L.seg) + lo;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x)
        seg[i].group_id = L.garbage[lo + i] ? -1 : L.inlier[lo + i];
}

} // namespace

void launch_filter_model(const Comps & K, const Lines & L, FrameCtrl * ctrl, int n_slots, float min_length, int rounds_total, bool refined, cudaStream_t st)
{
    LRB_CARVEOUT(k_filter_model); LRB_CARVEOUT(k_score); LRB_CARVEOUT(k_refit_assign); LRB_CARVEOUT(k_finalize);
    k_filter_model<<<n_slots, 1024, 0, st>>>(K, L, ctrl, min_length, rounds_total, refined ? refine_output(K, 0) : nullptr, (size_t)K.cap_pool);
}

void launch_ransac(const Lines & L, FrameCtrl * ctrl, int n_slots, const unsigned * stream, unsigned long long stream_len,
                   int T, cudaStream_t st, cudaEvent_t * ev_score)
{
    // 24 KB of shared memory per scoring CTA: state the split (four CTAs per SM) instead of leaving the SM at the driver's maximum (see launch_nms)
    static PerDeviceOnce once_cv;
    once_cv.run([&] { cudaFuncSetAttribute(k_score_wide, cudaFuncAttributePreferredSharedMemoryCarveout, 50); });
    for (int k = 0; k < kMaxModels; ++k) {
        launch_sampler(L, ctrl, n_slots, stream, stream_len, k, T, st);
        if (ev_score) cudaEventRecord(ev_score[2 * k], st);
        // a large group fills the GPU with one warp per 32 hypotheses; one or two frames need the observations split over warps too
        if (n_slots <= 2) k_score_wide<<<dim3((T + 31) / 32, n_slots), SCW_THREADS, 0, st>>>(L, ctrl, k, T);
        else k_score<<<dim3((T + SC_THREADS - 1) / SC_THREADS, n_slots), SC_THREADS, 0, st>>>(L, ctrl, k, T);
        if (ev_score) cudaEventRecord(ev_score[2 * k + 1], st);
        k_refit_assign<<<n_slots, 1024, 0, st>>>(L, ctrl, k, T);
    }
    dim3 gf(8, n_slots);
    k_finalize<<<gf, 256, 0, st>>>(L, ctrl);
}

} // namespace lrb
