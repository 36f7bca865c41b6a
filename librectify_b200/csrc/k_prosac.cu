// k_prosac.cu — the reference's dormant PROSAC estimator: PROSAC_Estimator<LinePencilModel>::solve,
// reference src/prosac.h:104-299 (growth function and stopping rule of Chum & Matas; never instantiated by the reference's build,
// pinned here against a harness instantiation of the compiled reference: oracle/ref_harness.cpp, tests/golden/prosac.npz).
//
// Input: the pencil model of all lines, the listed observations and their Hough weights (k_hough.cu computes both).  The
// estimator sorts the observations by weight (argsort = std::stable_sort descending, src/utils.h:36-44), then draws at most
// T_N = niter_RANSAC(0.9, 0.5, 2) = 9 semi-random pairs from the growing prefix U_n of that order, scores each against ALL
// observations and keeps the best; the stopping rule walks the inlier flags once per improvement.  Finally fit_optimal on
// the best hypothesis' inliers.  Quirk kept: the semi-random sample uses idx(n) for u_n (src/prosac.h:175), one past the paper's.
//
// GPU form: one CTA.  The stable order is a rank-by-counting (m^2 comparisons over 1024 threads), every hypothesis is scored by
// the whole CTA (error per observation, block-wide count), the O(9) control flow and the O(N) stopping-rule walk run on thread 0
// in double precision like the reference (same expressions, no contraction), the uniforms come from the engine's stream
// (mt_stream.cu: the generator seeded as the reference's, k = 2^32 u as integers; choice_knuth's test is exact in that form).
#include <climits>
#include <cstdlib>
#include "common.cuh"
#include "eig.cuh"
#include "kernels.h"
#include "pencil.cuh"

namespace lrb {

namespace {

constexpr int PS_THREADS = 1024;

struct ProsacOut { int iterations, n_final, n_star, I_N_best; int samples[9][2]; int I_N[9]; float vp[3]; int draws; };

__device__ int ps_niter(double p, double epsilon, int s, int Nmax)             // prosac.h:30-54
{
    if (Nmax == -1) Nmax = INT_MAX;
    if (epsilon <= 0.) return 1;
    const double logarg = -exp(__dmul_rn((double)s, log(__dsub_rn(1., epsilon))));
    const double logval = log(__dadd_rn(1., logarg));
    const double N = __ddiv_rn(log(__dsub_rn(1., p)), logval);
    if (logval < 0. && N < (double)Nmax) return (int)ceil(N);
    return Nmax;
}

// choice_knuth (math_utils.cpp:14-39) for a subset of size n out of N, on the integer stream: select <=> (N - t) k < (n - m) 2^32
__device__ void ps_choice(int N, int n, const unsigned * stream, unsigned long long & pos, int * out)
{
    int t = 0, m = 0;
    while (m < n) {
        const unsigned long long k = stream[pos++];
        if ((unsigned long long)(N - t) * k >= ((unsigned long long)(n - m) << 32)) { t++; }
        else { out[m] = t; t++; m++; }
    }
}

__global__ void __launch_bounds__(PS_THREADS)
k_prosac(const float * __restrict__ h, const float * __restrict__ anchor, const float * __restrict__ dir, const float * __restrict__ len,
         int n_lines, const int * __restrict__ idx, int N, const float * __restrict__ weights, const unsigned * __restrict__ stream,
         unsigned long long stream_len, float tol, int * sidx, unsigned char * flags /* 2 x N */, int * inl, ProsacOut * out, int dbg_stage)
{
    __shared__ int wsum[32];
    __shared__ float s_cov[9];
    __shared__ int s_state[8];           // 0: continue, 1: sample ok, 2: a, 3: b, 4: count, 5: which flag buffer holds the best, 6: have best
    __shared__ float s_p[3];
    const int t_ = threadIdx.x, lane = t_ & 31, warp = t_ >> 5;
    // ---- argsort(weights): stable, descending --------------------------------------------------------------------------------
    for (int i = t_; i < N; i += PS_THREADS) {
        const float wi = weights[i];
        int pos = 0;
        // "equal" as the sort sees it: neither is greater.  All weights are NaN when the winning Hough cell lies outside the unit
        // disc (sqrt of a negative number, line_pencil.cpp:83): every comparison is false and the stable sort keeps the list order.
        for (int j = 0; j < N; ++j) { const float wj = weights[j]; pos += (wj > wi) || (!(wi > wj) && j < i); }
        sidx[pos] = idx[i];
    }
    __syncthreads();
    if (dbg_stage == 1) return;
    // ---- state of the sampling loop (thread 0; prosac.h:62-66, 114-150) ------------------------------------------------------
    const float eta = 0.05f, beta = 0.01f, p_good_sample = 0.9f, max_outlier_proportion = 0.5f;
    const float chi2_value = 4.21788459f;                       // chi2_table[floor(clip(2 * 0.02f, 0.01f, 0.2f) * 100)] = chi2_table[4]
    const int m = 2;
    int T_N = 0, n_star = N, I_n_star = 0, I_N_best = 0, t = 0, n = m, T_n_prime = 1, k_n_star = 0, I_N_min = 0;
    double T_n = 0.;
    unsigned long long pos = 0;
    if (t_ == 0) {
        T_N = ps_niter((double)p_good_sample, (double)max_outlier_proportion, m, -1);
        k_n_star = T_N;
        I_N_min = (int)__dmul_rn(__dsub_rn(1., (double)max_outlier_proportion), (double)N);
        T_n = (double)T_N;
        for (int i = 0; i < m; i++) T_n = __dmul_rn(T_n, __ddiv_rn((double)(n - i), (double)(N - i)));
        s_state[5] = 0; s_state[6] = 0;
        out->iterations = 0;
        for (int i = 0; i < 9; ++i) { out->samples[i][0] = 0; out->samples[i][1] = 0; out->I_N[i] = 0; }
    }
    int cur = 0;                                                // flag buffer scored into next (uniform: mirrors s_state[5])
    while (true) {
        if (t_ == 0) {
            int go = (((I_N_best < I_N_min) || t <= k_n_star) && t < T_N) ? 1 : 0;
            if (go && pos + 64 > stream_len) go = 0;            // cannot happen: at most 9 samples of at most 12 draws
            s_state[0] = go; s_state[1] = 0;
            if (go) {
                t = t + 1;
                if ((t > T_n_prime) && (n < n_star)) {          // growth function g(t), prosac.h:158-169
                    const double T_nplus1 = __ddiv_rn(__dmul_rn(T_n, (double)(n + 1)), (double)(n + 1 - m));
                    n = n + 1;
                    T_n_prime = (int)((double)T_n_prime + ceil(__dsub_rn(T_nplus1, T_n)));
                    T_n = T_nplus1;
                }
                int sample[2];
                if (t > T_n_prime) ps_choice(n, m, stream, pos, sample);
                else { ps_choice(n - 1, m - 1, stream, pos, sample); sample[m - 1] = n; }
                out->samples[t - 1][0] = sample[0]; out->samples[t - 1][1] = sample[1]; out->I_N[t - 1] = -1; out->iterations = t;
                const int a = sidx[sample[0]], b = sidx[sample[1]];
                const float * A = h + 3 * a; const float * B = h + 3 * b;
                const float ex = __fsub_rn(A[0], B[0]), ey = __fsub_rn(A[1], B[1]), ez = __fsub_rn(A[2], B[2]);
                const bool ok = __fsqrt_rn(sum3(__fmul_rn(ex, ex), __fmul_rn(ey, ey), __fmul_rn(ez, ez))) > kDegeneracyTol;   // line_pencil.cpp:89-98
                float p0, p1, p2; cross3(A, B, p0, p1, p2);                                                                 // line_pencil.cpp:101-108
                s_p[0] = p0; s_p[1] = p1; s_p[2] = p2;
                s_state[1] = ok ? 1 : 0;
            }
        }
        __syncthreads();
        if (!s_state[0]) break;
        if (!s_state[1]) { __syncthreads(); continue; }         // sample_check failed: the draws are spent, nothing is scored
        // ---- support of the hypothesis over all observations (prosac.h:196-199) ---------------------------------------------
        const Hyp hp = make_hyp(s_p[0], s_p[1], s_p[2]);
        unsigned char * fl = flags + (size_t)cur * N;
        int cnt = 0;
        for (int k = t_; k < N; k += PS_THREADS) {
            const int q = sidx[k];
            const int in = line_error(hp, anchor[2 * q], anchor[2 * q + 1], dir[2 * q], dir[2 * q + 1]) < tol;
            fl[k] = (unsigned char)in; cnt += in;
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
        if (lane == 0) wsum[warp] = cnt;
        __syncthreads();
        if (t_ == 0) {
            int I_N = 0;
            for (int w = 0; w < PS_THREADS / 32; ++w) I_N += wsum[w];
            out->I_N[t - 1] = I_N;
            s_state[4] = 0;
            if (I_N > I_N_best) {                               // prosac.h:203-281
                I_N_best = I_N;
                s_state[4] = 1;                                 // the buffer just written becomes the best
                int n_best = N, I_n_best = I_N;
                double epsilon_n_best = __ddiv_rn((double)I_n_best, (double)n_best);
                for (int n_test = N, I_n_test = I_N; n_test > m; n_test--) {
                    if ((I_n_test * n_best > I_n_best * n_test) &&
                        ((double)I_n_test > __dadd_rn(__dmul_rn(epsilon_n_best, (double)n_test),
                                                      sqrt(__dmul_rn(__dmul_rn(__dmul_rn((double)n_test, epsilon_n_best), __dsub_rn(1., epsilon_n_best)), 2.706))))) {
                        // Imin(m, n_test), prosac.h:78-83
                        const double mu = (double)__fmul_rn((float)n_test, beta);
                        const double sigma = sqrt((double)__fmul_rn(__fmul_rn((float)n_test, beta), __fsub_rn(1.0f, beta)));
                        const int imin = (int)ceil(__dadd_rn(__dadd_rn((double)m, mu), __dmul_rn(sigma, sqrt((double)chi2_value))));
                        if (I_n_test < imin) break;
                        n_best = n_test; I_n_best = I_n_test; epsilon_n_best = __ddiv_rn((double)I_n_best, (double)n_best);
                    }
                    I_n_test -= fl[n_test - 1];
                }
                if (I_n_best * n_star > I_n_star * n_best) {
                    n_star = n_best; I_n_star = I_n_best;
                    k_n_star = ps_niter(__dsub_rn(1., (double)eta), __dsub_rn(1., __ddiv_rn((double)I_n_star, (double)n_star)), m, T_N);
                }
            }
        }
        __syncthreads();
        if (s_state[4]) { if (t_ == 0) { s_state[5] = cur; s_state[6] = 1; } cur ^= 1; }
        __syncthreads();
    }
    if (dbg_stage == 2) return;
    // ---- fit_optimal on the inliers of the best hypothesis, in sorted order (prosac.h:296-297) -----------------------------------
    const bool have = s_state[6] != 0;
    const unsigned char * bf = flags + (size_t)s_state[5] * N;
    int ni = 0;
    for (int k0 = 0; k0 < N; k0 += PS_THREADS) {
        const int k = k0 + t_;
        const int f = (have && k < N) ? bf[k] : 0;
        int total; const int p = ni + block_scan_flag(f, wsum, total);
        if (f) inl[p] = sidx[k];
        ni += total;
    }
    __syncthreads();
    const bool all = (ni == 0);                                 // empty list (nothing was ever scored) => all lines (line_pencil.cpp:114-117)
    const int cntl = all ? n_lines : ni;
    if (t_ < 6) {
        const int r = (t_ == 0) ? 0 : (t_ < 3 ? 1 : 2);
        const int c = (t_ == 0) ? 0 : (t_ == 1 ? 0 : (t_ == 2 ? 1 : t_ - 3));
        float a = 0.0f;
        for (int kk = 0; kk < cntl; ++kk) {
            const int q = all ? kk : inl[kk];
            const float v = __fmul_rn(__fmul_rn(h[3 * q + r], len[q]), h[3 * q + c]);
            a = (kk == 0) ? v : __fadd_rn(a, v);
        }
        s_cov[r * 3 + c] = a; s_cov[c * 3 + r] = a;
    }
    __syncthreads();
    if (t_ == 0) {
        float eval[3], evec[9];
        sym_eig<3>(s_cov, eval, evec);
        int kk = 0;
        for (int i = 1; i < 3; ++i) if (eval[i] < eval[kk]) kk = i;
        out->vp[0] = evec[kk * 3]; out->vp[1] = evec[kk * 3 + 1]; out->vp[2] = evec[kk * 3 + 2];
        out->n_final = n; out->n_star = n_star; out->I_N_best = I_N_best; out->draws = (int)pos;
    }
}

} // namespace

size_t prosac_out_bytes() { return sizeof(ProsacOut); }

void launch_prosac(const float * h, const float * anchor, const float * dir, const float * len, int n_lines, const int * idx, int m,
                   const float * weights, const unsigned * stream, unsigned long long stream_len, float tol,
                   int * sidx, unsigned char * flags, int * inl, void * out, cudaStream_t st)
{
    static const int dbg = [] { const char * v = getenv("LRB_PROSAC_STAGE"); return v ? atoi(v) : 0; }();
    k_prosac<<<1, PS_THREADS, 0, st>>>(h, anchor, dir, len, n_lines, idx, m, weights, stream, stream_len, tol, sidx, flags, inl,
                                       static_cast<ProsacOut *>(out), dbg);
}

} // namespace lrb
