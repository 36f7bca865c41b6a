// k_hough.cu — the reference's Hough-voted estimator (dormant in the reference's build, reachable
// through its templates): LinePencilModel::get_weights (reference src/line_pencil.cpp:47-86) and DirectEstimator::solve
// (reference src/estimator.h:82-96).
//
//   get_weights: 20 000 hypotheses; each draws two listed lines with std::mt19937() (default seed 5489) +
//   std::uniform_int_distribution<int>(0, m-1), intersects them (h_a x h_b), normalises the point onto the upper unit
//   hemisphere and votes length(a) + length(b) into cell (round(31 x + 32), round(31 y + 32)) of a 65 x 65 accumulator;
//   the first maximum in column-major order gives the dominant vanishing direction p, and the weight of every listed
//   line is inclination(line, p)^4.  DirectEstimator: inliers = weights > 0.95f, then fit_optimal on them.
//
// GPU form: one CTA per problem.  The draws are resolved in parallel (the generator's raw outputs come from
// mt_stream.cu; libstdc++'s uniform_int_distribution is Lemire's multiply-shift with a rejection that is compacted by a
// block scan), every thread evaluates its hypotheses and votes with shared-memory atomics into a privatised 65 x 65
// accumulator.  Float addition does not commute with the reference's sequential `accumulator(u,v) +=`, so the atomic
// sums only SELECT candidates: every cell within 1 % of the largest atomic sum (the sequential and the atomic sum of a
// cell differ by at most count * 2^-24 relative <= 0.12 %) is re-summed by one warp in hypothesis order — bit-identical
// to the reference — and the first strict maximum of those exact sums in column-major order wins.
#include "common.cuh"
#include "eig.cuh"
#include "kernels.h"
#include "pencil.cuh"

namespace lrb {

namespace {

constexpr int kHtSpace = 65, kHtCells = kHtSpace * kHtSpace, kHtHyp = 20000;     // line_pencil.h:22-23
constexpr int kHtDraws = 2 * kHtHyp;
constexpr int HT_THREADS = 1024;

// LinePencilModel on ALL lines of the problem (reference line_pencil.cpp:25-32 after normalize_lines, geometry.cpp:96-112,258-282)
__global__ void __launch_bounds__(HT_THREADS)
k_hough_model(const SegRec * __restrict__ seg, int n, float * __restrict__ h, float * __restrict__ anchor, float * __restrict__ dir,
              float * __restrict__ len)
{
    __shared__ float red[4][32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    float mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY;
    for (int i = t; i < n; i += HT_THREADS) {
        const SegRec l = seg[i];
        mnx = fminf(mnx, fminf(l.x1, l.x2)); mxx = fmaxf(mxx, fmaxf(l.x1, l.x2));
        mny = fminf(mny, fminf(l.y1, l.y2)); mxy = fmaxf(mxy, fmaxf(l.y1, l.y2));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, s)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, s));
        mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, s)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, s));
    }
    if (lane == 0) { red[0][warp] = mnx; red[1][warp] = mny; red[2][warp] = mxx; red[3][warp] = mxy; }
    __syncthreads();
    mnx = red[0][0]; mny = red[1][0]; mxx = red[2][0]; mxy = red[3][0];
    for (int w = 1; w < HT_THREADS / 32; ++w) {
        mnx = fminf(mnx, red[0][w]); mny = fminf(mny, red[1][w]); mxx = fmaxf(mxx, red[2][w]); mxy = fmaxf(mxy, red[3][w]);
    }
    const float sx = __fsub_rn(mxx, mnx), sy = __fsub_rn(mxy, mny);
    const float cx = __fadd_rn(mnx, __fmul_rn(0.5f, sx)), cy = __fadd_rn(mny, __fmul_rn(0.5f, sy));      // bbox_center
    const float scale = fmaxf(sx, sy);                                                                    // bbox_size().maxCoeff()
    for (int i = t; i < n; i += HT_THREADS) {
        const SegRec l = seg[i];
        const float x1 = __fdiv_rn(__fsub_rn(l.x1, cx), scale), y1 = __fdiv_rn(__fsub_rn(l.y1, cy), scale);
        const float x2 = __fdiv_rn(__fsub_rn(l.x2, cx), scale), y2 = __fdiv_rn(__fsub_rn(l.y2, cy), scale);
        const float hx = __fsub_rn(y1, y2), hy = __fsub_rn(x2, x1), hz = __fsub_rn(__fmul_rn(x1, y2), __fmul_rn(y1, x2));   // (x1,y1,1) x (x2,y2,1)
        const float nrm = __fsqrt_rn(sum3(__fmul_rn(hx, hx), __fmul_rn(hy, hy), __fmul_rn(hz, hz)));
        h[3 * i] = __fdiv_rn(hx, nrm); h[3 * i + 1] = __fdiv_rn(hy, nrm); h[3 * i + 2] = __fdiv_rn(hz, nrm);
        anchor[2 * i] = __fdiv_rn(__fadd_rn(x2, x1), 2.0f); anchor[2 * i + 1] = __fdiv_rn(__fadd_rn(y2, y1), 2.0f);
        const float ddx = __fsub_rn(x2, x1), ddy = __fsub_rn(y2, y1);
        const float ln = __fsqrt_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)));
        len[i] = ln;
        dir[2 * i] = __fdiv_rn(ddx, ln); dir[2 * i + 1] = __fdiv_rn(ddy, ln);
    }
}

struct HoughOut { int cell_u, cell_v, n_inliers, n_candidates; float p[3]; float vp[3]; };

__global__ void __launch_bounds__(HT_THREADS)
k_hough_direct(const float * __restrict__ h, const float * __restrict__ anchor, const float * __restrict__ dir, const float * __restrict__ len,
               int n_lines, const int * __restrict__ idx, int m, const unsigned * __restrict__ raw, int n_raw,
               int * __restrict__ draws /* kHtDraws */, int * __restrict__ hyp_cell /* kHtHyp */, float * __restrict__ hyp_val /* kHtHyp */,
               float * __restrict__ weights /* m */, int * __restrict__ inl /* m */, HoughOut * __restrict__ out)
{
    __shared__ float acc[kHtCells];
    __shared__ float exact[kHtCells];          // exact (sequential) sums of the candidate cells, -1 elsewhere
    __shared__ int wsum[32];
    __shared__ int s_n;
    __shared__ unsigned s_maxbits;
    __shared__ float s_p[3], s_cov[9], s_vp[3];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int c = t; c < kHtCells; c += HT_THREADS) { acc[c] = 0.0f; exact[c] = -1.0f; }
    if (t == 0) { s_n = 0; s_maxbits = 0u; }
    __syncthreads();

    // ---- the draws: libstdc++ uniform_int_distribution<int>(0, m-1) over mt19937 (Lemire, 64-bit product) -------------
    const unsigned um = (unsigned)m;
    const unsigned threshold = (0u - um) % um;
    int have = 0;                                                           // accepted draws so far
    for (int base = 0; base < n_raw && have < kHtDraws; base += HT_THREADS) {
        const int j = base + t;
        unsigned long long prod = 0ull; int ok = 0;
        if (j < n_raw) { prod = (unsigned long long)raw[j] * um; ok = ((unsigned)prod >= threshold); }
        int total; const int pos = have + block_scan_flag(ok, wsum, total);
        if (ok && pos < kHtDraws) draws[pos] = (int)(prod >> 32);
        have += total;
    }
    __syncthreads();
    if (have < kHtDraws) { if (t == 0) out->n_candidates = -1; return; }   // stream too short (the host sizes it with slack)

    // ---- hypotheses and votes ------------------------------------------------------------------------------------------
    const float k = 32.0f, k1 = 31.0f;                                      // floor(65 / 2.f), k - 1
    for (int i = t; i < kHtHyp; i += HT_THREADS) {
        const int a = idx[draws[2 * i]], b = idx[draws[2 * i + 1]];
        float x0, x1, x2;
        cross3(h + 3 * a, h + 3 * b, x0, x1, x2);
        int cell = -1; float val = 0.0f;
        if (!(fabsf(x0) < 0.0001f && fabsf(x1) < 0.0001f && fabsf(x2) < 0.0001f)) {
            const float z = sum3(__fmul_rn(x0, x0), __fmul_rn(x1, x1), __fmul_rn(x2, x2));
            if (z > 0.0f) { const float nr = __fsqrt_rn(z); x0 = __fdiv_rn(x0, nr); x1 = __fdiv_rn(x1, nr); x2 = __fdiv_rn(x2, nr); }
            if (x2 < 0.0f) { x0 = -x0; x1 = -x1; }
            const int u = (int)roundf(__fadd_rn(__fmul_rn(k1, x0), k)), v = (int)roundf(__fadd_rn(__fmul_rn(k1, x1), k));
            cell = u + kHtSpace * v;                                        // column-major, like the reference's ArrayXXf
            val = __fadd_rn(len[a], len[b]);
            atomicAdd(&acc[cell], val);
        }
        hyp_cell[i] = cell; hyp_val[i] = val;
    }
    __syncthreads();
    // ---- candidates: every cell within 1 % of the largest atomic sum gets its exact sequential sum --------------------
    unsigned mb = 0u;
    for (int c = t; c < kHtCells; c += HT_THREADS) mb = max(mb, __float_as_uint(acc[c]));   // sums are >= 0: bit order = value order
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) mb = max(mb, __shfl_xor_sync(0xffffffffu, mb, s));
    if (lane == 0) atomicMax(&s_maxbits, mb);
    __syncthreads();
    const float cut = __fmul_rn(__uint_as_float(s_maxbits), 0.99f);
    for (int c0 = 0; c0 < kHtCells; c0 += HT_THREADS / 32) {               // one warp per cell, candidates are rare
        const int c = c0 + warp;
        if (c < kHtCells && acc[c] >= cut && acc[c] > 0.0f) {
            float sum = 0.0f;
            for (int i0 = 0; i0 < kHtHyp; i0 += 32) {
                const int i = i0 + lane;
                const bool hit = i < kHtHyp && hyp_cell[i] == c;
                const float v = hit ? hyp_val[i] : 0.0f;
                unsigned mm = __ballot_sync(0xffffffffu, hit);
                while (mm) {                                                // in hypothesis order, as `accumulator(u,v) +=` runs
                    const int src = __ffs(mm) - 1; mm &= mm - 1;
                    sum = __fadd_rn(sum, __shfl_sync(0xffffffffu, v, src));
                }
            }
            if (lane == 0) { exact[c] = sum; atomicAdd(&s_n, 1); }
        }
    }
    __syncthreads();
    if (t == 0) {
        // maxCoeff(&u, &v): first strict maximum, column by column; all other cells are below every candidate
        int best = 0; float bv = exact[0];
        for (int c = 1; c < kHtCells; ++c) if (exact[c] > bv) { bv = exact[c]; best = c; }
        if (bv < 0.0f) best = 0;                                            // no vote at all: the reference's argmax of an all-zero array
        const int mu = best % kHtSpace, mv = best / kHtSpace;
        float p0 = __fdiv_rn(__fsub_rn((float)mu, k), k1), p1 = __fdiv_rn(__fsub_rn((float)mv, k), k1), p2 = 0.0f;
        const float nrm = __fsqrt_rn(sum3(__fmul_rn(p0, p0), __fmul_rn(p1, p1), 0.0f));
        if (nrm > 1.0f) { p0 = __fdiv_rn(p0, nrm); p1 = __fdiv_rn(p1, nrm); p2 = __fdiv_rn(p2, nrm); }
        p2 = __fsqrt_rn(__fsub_rn(1.0f, __fadd_rn(__fmul_rn(p0, p0), __fmul_rn(p1, p1))));     // pow(x, 2.f) == x * x, correctly rounded
        s_p[0] = p0; s_p[1] = p1; s_p[2] = p2;
        out->cell_u = mu; out->cell_v = mv; out->n_candidates = s_n; out->p[0] = p0; out->p[1] = p1; out->p[2] = p2;
    }
    __syncthreads();
    // ---- weights = inclination^4 (pow(float, int) promotes to double), inliers in list order -----------------------------
    const Hyp hp = make_hyp(s_p[0], s_p[1], s_p[2]);
    int ni = 0;
    for (int q0 = 0; q0 < m; q0 += HT_THREADS) {
        const int qi = q0 + t; int f = 0, q = 0;
        if (qi < m) {
            q = idx[qi];
            // inclination(), geometry.cpp:214-229 (line_error in pencil.cuh returns 1 - inclination; the weight needs the value itself)
            float vx, vy;
            if (hp.ideal) { vx = hp.px; vy = hp.py; }
            else { vx = __fadd_rn(-anchor[2 * q], hp.px); vy = __fadd_rn(-anchor[2 * q + 1], hp.py); }
            const float z = __fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy));
            if (z > 0.0f) { const float nr = __fsqrt_rn(z); vx = __fdiv_rn(vx, nr); vy = __fdiv_rn(vy, nr); }
            const float inc = fabsf(__fadd_rn(__fmul_rn(vx, dir[2 * q]), __fmul_rn(vy, dir[2 * q + 1])));
            const double d = (double)inc, d2 = __dmul_rn(d, d);
            const float w = __double2float_rn(__dmul_rn(d2, d2));
            weights[qi] = w;
            f = w > 0.95f;
        }
        int total; const int pos = ni + block_scan_flag(f, wsum, total);
        if (f) inl[pos] = q;
        ni += total;
    }
    __syncthreads();
    // ---- fit_optimal (line_pencil.cpp:111-128): empty inlier list => all lines of the model --------------------------------
    const bool all = (ni == 0);
    const int cnt = all ? n_lines : ni;
    if (t < 6) {
        const int r = (t == 0) ? 0 : (t < 3 ? 1 : 2);
        const int c = (t == 0) ? 0 : (t == 1 ? 0 : (t == 2 ? 1 : t - 3));
        float a = 0.0f;
        for (int kk = 0; kk < cnt; ++kk) {
            const int q = all ? kk : inl[kk];
            const float v = __fmul_rn(__fmul_rn(h[3 * q + r], len[q]), h[3 * q + c]);
            a = (kk == 0) ? v : __fadd_rn(a, v);
        }
        s_cov[r * 3 + c] = a; s_cov[c * 3 + r] = a;
    }
    __syncthreads();
    if (t == 0) {
        float eval[3], evec[9];
        sym_eig<3>(s_cov, eval, evec);
        int kk = 0;
        for (int i = 1; i < 3; ++i) if (eval[i] < eval[kk]) kk = i;
        out->vp[0] = evec[kk * 3]; out->vp[1] = evec[kk * 3 + 1]; out->vp[2] = evec[kk * 3 + 2];
        out->n_inliers = ni;
    }
}

} // namespace

size_t hough_out_bytes() { return sizeof(HoughOut); }

void launch_hough_direct(const float * seg7, int n_lines, const int * idx, int m, const unsigned * raw, int n_raw,
                         float * h, float * anchor, float * dir, float * len, int * draws, int * hyp_cell, float * hyp_val,
                         float * weights, int * inl, void * out, cudaStream_t st)
{
    k_hough_model<<<1, HT_THREADS, 0, st>>>(reinterpret_cast<const SegRec *>(seg7), n_lines, h, anchor, dir, len);
    k_hough_direct<<<1, HT_THREADS, 0, st>>>(h, anchor, dir, len, n_lines, idx, m, raw, n_raw, draws, hyp_cell, hyp_val, weights, inl,
                                             static_cast<HoughOut *>(out));
}

int hough_draws() { return kHtDraws; }
int hough_hypotheses() { return kHtHyp; }

} // namespace lrb
