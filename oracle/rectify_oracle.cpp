/* TEST INFRASTRUCTURE ONLY — CPU restatement of librectify's hot path (see rectify_oracle.h).
 *
 * Plain C++14, no Eigen, no OpenMP, compiled strict-IEEE (-O2 -fno-fast-math -ffp-contract=off) so that every float
 * expression below rounds exactly as written.  "ref" below means /root/reference/src of RomanJuranek/librectify; the
 * arithmetic orders are those of the reference compiled in its canonical form (scalar Eigen, no FMA — oracle/Makefile,
 * librectify_canon.so), against which every stage here is pinned by tests/test_oracle_vs_ref.py.
 */
#include "rectify_oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <deque>
#include <limits>
#include <map>
#include <set>
#include <vector>

namespace {

using std::vector;
typedef OrcLineSegment Seg;

/* ---- constants of ref config.h --------------------------------------------------------------------------------- */
const float kEps = 1e-6f;                 /* config.h:10 */
const int   kMaxModels = 4;               /* config.h:23 */
const float kInlierDeg = 2.0f;            /* config.h:25 */
const float kGarbageDeg = 4.0f;           /* config.h:27 */
const int   kRansacIter = 10000;          /* config.h:29 */
const float kSeedRatio = 0.95f;           /* config.h:44 (double literal narrowed at the float parameter) */
const float kTraceTol = 0.25f;            /* config.h:47 */
const float kLineMaxErr = 2.0f;           /* config.h:50 */
const float kLineMinLength = 5.f;         /* config.h:56 */
const int   kComponentMinSize = 5;        /* config.h:59 */
const float kDegeneracyTol = 0.05f;       /* line_pencil.h:21 */

unsigned g_seed = 0;

/* ---- a1: image_from_buffer (ref image.cpp:11-20) ------------------------------------------------------------------ */
vector<float> image_from_buffer(const float * buffer, int w, int h, int stride)
{
    if (stride < 0) { buffer = buffer + (long long)(h - 1) * stride; stride = -stride; }
    vector<float> im(size_t(w) * h);
    for (int r = 0; r < h; ++r) std::memcpy(&im[size_t(r) * w], buffer + (long long)r * stride, sizeof(float) * w);
    return im;
}

/* ---- a2: gauss_deriv_kernel (ref filter.cpp:65-78), size=2, sigma=1 ----------------------------------------------- */
void gauss_deriv_kernel(bool dir_x, float H[25])
{
    const int size = 2; const float sigma = 1.0f;
    for (int i = 0; i < 5; ++i)
        for (int j = 0; j < 5; ++j) {
            float x = float(j - size), y = float(i - size);
            float z = dir_x ? x : y;
            H[i*5+j] = z / float(2*M_PI*std::pow(sigma, 4.0f)) * std::exp(-(std::pow(x,2.0f)+std::pow(y,2.0f))/(2*std::pow(sigma,2.0f)));
        }
}

/* ---- a3: conv_2d (ref filter.cpp:81-98): direct correlation, row-major sequential sum, 2-px border zero ------------ */
void conv5(const vector<float> & im, int w, int h, const float K[25], vector<float> & out)
{
    out.assign(size_t(w) * h, 0.f);
    for (int i = 0; i + 5 <= h; ++i)
        for (int j = 0; j + 5 <= w; ++j) {
            float acc = im[size_t(i) * w + j] * K[0];
            for (int t = 1; t < 25; ++t) {
                int a = t / 5, b = t % 5;
                acc = acc + im[size_t(i + a) * w + (j + b)] * K[t];
            }
            out[size_t(i + 2) * w + (j + 2)] = acc;
        }
}

/* ---- a5: orientation tables (ref line_detector.cpp:142-146): theta=float(i*M_PI)/8, sinf/cosf ---------------------- */
void orientation_tables(float s[8], float c[8])
{
    for (int i = 0; i < 8; ++i) { float theta = float(i * M_PI) / 8; s[i] = std::sin(theta); c[i] = std::cos(theta); }
}

struct Stages {
    int w, h;
    vector<float> dx, dy, mag;
    vector<int> bin;
    vector<float> grad[8];
    vector<int> seed_r, seed_c;
    vector<vector<int> > comp_px;     /* raster index, BFS order, kept components only */
    vector<vector<float> > comp_val;
    vector<Seg> lines;
};

/* ---- a10: SelfAdjointEigenSolver<Matrix2f/3f>::compute (Eigen SelfAdjointEigenSolver.h:407-453,490-551,817-870,
 *           Jacobi.h:231-268, Tridiagonalization.h:464-504) — column-major Q ---------------------------------------- */
struct Givens { float c, s; };
Givens make_givens(float p, float q)
{
    Givens g;
    if (q == 0.f) { g.c = p < 0.f ? -1.f : 1.f; g.s = 0.f; }
    else if (p == 0.f) { g.c = 0.f; g.s = q < 0.f ? 1.f : -1.f; }
    else if (std::fabs(p) > std::fabs(q)) {
        float t = q / p; float u = std::sqrt(1.f + t * t); if (p < 0.f) u = -u;
        g.c = 1.f / u; g.s = -t * g.c;
    } else {
        float t = p / q; float u = std::sqrt(1.f + t * t); if (q < 0.f) u = -u;
        g.s = -1.f / u; g.c = -t * g.s;
    }
    return g;
}

float eig_hypot(float x, float y)
{
    x = std::fabs(x); y = std::fabs(y);
    float p = std::max(x, y);
    if (p == 0.f) return 0.f;
    float qp = std::min(y, x) / p;
    return p * std::sqrt(1.f + qp * qp);
}

/* Q is n x n, Q[col*n + row] */
void tridiagonal_qr_step(float * diag, float * subdiag, int start, int end, float * Q, int n)
{
    float td = (diag[end-1] - diag[end]) * 0.5f;
    float e = subdiag[end-1];
    float mu = diag[end];
    if (td == 0.f) mu -= std::fabs(e);
    else {
        float e2 = e * e;
        float h = eig_hypot(td, e);
        if (e2 == 0.f) mu -= (e / (td + (td > 0.f ? 1.f : -1.f))) * (e / h);
        else           mu -= e2 / (td + (td > 0.f ? h : -h));
    }
    float x = diag[start] - mu;
    float z = subdiag[start];
    for (int k = start; k < end; ++k) {
        Givens rot = make_givens(x, z);
        float sdk  = rot.s * diag[k] + rot.c * subdiag[k];
        float dkp1 = rot.s * subdiag[k] + rot.c * diag[k+1];
        diag[k]   = rot.c * (rot.c * diag[k] - rot.s * subdiag[k]) - rot.s * (rot.c * subdiag[k] - rot.s * diag[k+1]);
        diag[k+1] = rot.s * sdk + rot.c * dkp1;
        subdiag[k] = rot.c * sdk - rot.s * dkp1;
        if (k > start) subdiag[k-1] = rot.c * subdiag[k-1] - rot.s * z;
        x = subdiag[k];
        if (k < end - 1) { z = -rot.s * subdiag[k+1]; subdiag[k+1] = rot.c * subdiag[k+1]; }
        /* Q = Q * G : applyOnTheRight(k,k+1,rot) with j.transpose() -> (c,-s) */
        if (!(rot.c == 1.f && rot.s == 0.f))
            for (int i = 0; i < n; ++i) {
                float xi = Q[k*n+i], yi = Q[(k+1)*n+i];
                Q[k*n+i]     = rot.c * xi - rot.s * yi;
                Q[(k+1)*n+i] = rot.s * xi + rot.c * yi;
            }
    }
}

/* A: symmetric n x n (only lower triangle read), A[row*n+col]. eval ascending, evec column-major evec[col*n+row] */
void self_adjoint_eig(const float * A, int n, float * eval, float * evec)
{
    float m[9] = {0};   /* m[row*n+col], lower triangle copy */
    for (int r = 0; r < n; ++r) for (int c = 0; c <= r; ++c) m[r*n+c] = A[r*n+c];
    float scale = 0.f;
    for (int i = 0; i < n*n; ++i) scale = std::max(scale, std::fabs(m[i]));
    if (scale == 0.f) scale = 1.f;
    for (int r = 0; r < n; ++r) for (int c = 0; c <= r; ++c) m[r*n+c] /= scale;
    float diag[3], sub[2] = {0, 0};
    float Q[9] = {0};
    if (n == 2) {
        /* generic Householder path degenerates to identity for n=2 (tail of length 0 => tau=0, beta=a10) */
        diag[0] = m[0]; diag[1] = m[3]; sub[0] = m[2];
        Q[0] = 1.f; Q[3] = 1.f;
    } else {
        const float tol = std::numeric_limits<float>::min();
        diag[0] = m[0];
        float v1norm2 = m[6] * m[6];
        if (v1norm2 <= tol) {
            diag[1] = m[4]; diag[2] = m[8]; sub[0] = m[3]; sub[1] = m[7];
            Q[0] = Q[4] = Q[8] = 1.f;
        } else {
            float beta = std::sqrt(m[3] * m[3] + v1norm2);
            float invBeta = 1.f / beta;
            float m01 = m[3] * invBeta, m02 = m[6] * invBeta;
            float q = 2.f * m01 * m[7] + m02 * (m[8] - m[4]);
            diag[1] = m[4] + m02 * q;
            diag[2] = m[8] - m02 * q;
            sub[0] = beta;
            sub[1] = m[7] - m01 * q;
            /* mat << 1,0,0, 0,m01,m02, 0,m02,-m01 (row-wise fill) -> column-major Q */
            Q[0] = 1.f; Q[1*3+1] = m01; Q[2*3+1] = m02; Q[1*3+2] = m02; Q[2*3+2] = -m01;
        }
    }
    /* computeFromTridiagonal_impl */
    int end = n - 1, start = 0, iter = 0;
    const int maxIter = 30;
    const float considerAsZero = std::numeric_limits<float>::min();
    const float precision = 2.f * std::numeric_limits<float>::epsilon();
    while (end > 0) {
        for (int i = start; i < end; ++i)
            if (std::fabs(sub[i]) <= (std::fabs(diag[i]) + std::fabs(diag[i+1])) * precision || std::fabs(sub[i]) <= considerAsZero)
                sub[i] = 0.f;
        while (end > 0 && sub[end-1] == 0.f) end--;
        if (end <= 0) break;
        iter++;
        if (iter > maxIter * n) break;
        start = end - 1;
        while (start > 0 && sub[start-1] != 0.f) start--;
        tridiagonal_qr_step(diag, sub, start, end, Q, n);
    }
    if (iter <= maxIter * n) {
        for (int i = 0; i < n - 1; ++i) {
            int k = 0;
            for (int j = 1; j < n - i; ++j) if (diag[i+j] < diag[i+k]) k = j;   /* minCoeff: first minimum */
            if (k > 0) {
                std::swap(diag[i], diag[k+i]);
                for (int r = 0; r < n; ++r) std::swap(Q[i*n+r], Q[(k+i)*n+r]);
            }
        }
    }
    for (int i = 0; i < n; ++i) eval[i] = diag[i] * scale;
    for (int i = 0; i < n*n; ++i) evec[i] = Q[i];
}

/* ---- a10: fit_line_parameters (ref geometry.cpp:20-61); X rows are (row, col) ------------------------------------ */
Seg fit_line_parameters(const vector<float> & Xr, const vector<float> & Xc, const vector<float> & w)
{
    size_t n = w.size();
    float wsum = w[0];
    for (size_t i = 1; i < n; ++i) wsum = wsum + w[i];
    vector<float> wt(n);
    for (size_t i = 0; i < n; ++i) wt[i] = w[i] / wsum;
    float a0 = Xr[0] * wt[0], a1 = Xc[0] * wt[0];
    for (size_t i = 1; i < n; ++i) { a0 = a0 + Xr[i] * wt[i]; a1 = a1 + Xc[i] * wt[i]; }
    vector<float> c0(n), c1(n);
    for (size_t i = 0; i < n; ++i) { c0[i] = Xr[i] - a0; c1[i] = Xc[i] - a1; }
    /* cov = centered^T * diag(wt) * centered: (c(k,i)*wt(k)) * c(k,j), sequential in k */
    float cov[4];
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2; ++j) {
            const vector<float> & ci = i ? c1 : c0; const vector<float> & cj = j ? c1 : c0;
            float acc = (ci[0] * wt[0]) * cj[0];
            for (size_t k = 1; k < n; ++k) acc = acc + (ci[k] * wt[k]) * cj[k];
            cov[i*2+j] = acc;
        }
    float eval[2], evec[4];
    self_adjoint_eig(cov, 2, eval, evec);
    float d0 = evec[1*2+0], d1 = evec[1*2+1];   /* col(1): largest eigenvalue = direction */
    float n0 = evec[0*2+0], n1 = evec[0*2+1];   /* col(0): normal */
    float t0 = 0, t1 = 0, esum = 0;
    for (size_t k = 0; k < n; ++k) {
        float t = c0[k] * d0 + c1[k] * d1;
        if (k == 0) { t0 = t1 = t; } else { if (t < t0) t0 = t; if (t > t1) t1 = t; }
        float e = std::fabs(c0[k] * n0 + c1[k] * n1);
        esum = (k == 0) ? e : esum + e;
    }
    float wmean = wsum / float(n);
    Seg l;
    float p0r = a0 + d0 * t0, p0c = a1 + d1 * t0, p1r = a0 + d0 * t1, p1c = a1 + d1 * t1;
    l.x1 = p0c; l.y1 = p0r; l.x2 = p1c; l.y2 = p1r;
    l.weight = wmean;
    l.err = esum / float(n);
    l.group_id = -1;
    return l;
}

/* ---- a1..a10: find_line_segments (ref line_detector.cpp:185-251) ------------------------------------------------- */
void detect(const vector<float> & im, int w, int h, Stages & S, bool keep_planes)
{
    S.w = w; S.h = h;
    size_t npx = size_t(w) * h;
    float Hx[25], Hy[25];
    gauss_deriv_kernel(true, Hx); gauss_deriv_kernel(false, Hy);
    conv5(im, w, h, Hx, S.dx); conv5(im, w, h, Hy, S.dy);       /* image_gradients, line_detector.cpp:41-49 */
    S.mag.resize(npx);
    for (size_t i = 0; i < npx; ++i) S.mag[i] = std::sqrt(S.dx[i] * S.dx[i] + S.dy[i] * S.dy[i]);

    /* gradient_directions, line_detector.cpp:126-182 */
    float sn[8], cs[8];
    orientation_tables(sn, cs);
    for (int b = 0; b < 8; ++b) {
        S.grad[b].resize(npx);
        for (size_t i = 0; i < npx; ++i) S.grad[b][i] = std::fabs(S.dx[i] * sn[b] + S.dy[i] * cs[b]);
    }
    S.bin.assign(npx, 0);               /* uninitialised in the reference (line_detector.cpp:128); defined as 0 here */
    {
        vector<float> gmax(npx, 0.f);
        for (int b = 0; b < 8; ++b)
            for (size_t i = 0; i < npx; ++i) {
                if (S.grad[b][i] > gmax[i]) S.bin[i] = b;
                gmax[i] = std::max(gmax[i], S.grad[b][i]);
            }
    }
    for (int b = 0; b < 8; ++b) {       /* binary_dilate(grad_bin==b) then select, filter.cpp:46-62 */
        vector<unsigned char> mask(npx, 0);
        for (int r = 1; r + 1 < h; ++r)
            for (int c = 1; c + 1 < w; ++c) {
                bool any = false;
                for (int a = -1; a <= 1 && !any; ++a)
                    for (int d = -1; d <= 1; ++d)
                        if (S.bin[size_t(r + a) * w + (c + d)] == b) { any = true; break; }
                mask[size_t(r) * w + c] = any;
            }
        for (size_t i = 0; i < npx; ++i) if (!mask[i]) S.grad[b][i] = 0.f;
    }

    /* seeds: line_detector.cpp:209, find_peaks filter.cpp:161-195 */
    float maxmag = S.mag[0];
    for (size_t i = 1; i < npx; ++i) maxmag = std::max(maxmag, S.mag[i]);
    float min_seed = maxmag * (1 - std::max(std::min(kSeedRatio, 1.f), 0.f));
    struct Peak { int i, j; float v; };
    vector<Peak> peaks;
    for (int r = 0; r < h; ++r)
        for (int c = 0; c < w; ++c) {
            float mx = 0.f;                                   /* maximum_filter output is 0 on the 2-px border */
            if (r >= 2 && r + 2 < h && c >= 2 && c + 2 < w) {
                mx = S.mag[size_t(r - 2) * w + (c - 2)];
                for (int a = -2; a <= 2; ++a) for (int d = -2; d <= 2; ++d) mx = std::max(mx, S.mag[size_t(r + a) * w + (c + d)]);
            }
            float v = S.mag[size_t(r) * w + c];
            if (mx == v && v > min_seed) peaks.push_back(Peak{r, c, v});
        }
    std::sort(peaks.begin(), peaks.end(), [](const Peak & a, const Peak & b) { return a.v > b.v; });
    S.seed_r.resize(peaks.size()); S.seed_c.resize(peaks.size());
    for (size_t i = 0; i < peaks.size(); ++i) { S.seed_r[i] = peaks[i].i; S.seed_c[i] = peaks[i].j; }

    /* find_components line_detector.cpp:92-122 + flood filter.cpp:101-153 */
    vector<unsigned char> visited(npx, 0);
    for (int c = 0; c < w; ++c) { visited[c] = 1; visited[size_t(h - 1) * w + c] = 1; }
    for (int r = 0; r < h; ++r) { visited[size_t(r) * w] = 1; visited[size_t(r) * w + (w - 1)] = 1; }
    static const int dr[8] = {0, 0, 1, -1, -1, -1, 1, 1};
    static const int dc[8] = {-1, 1, 0, 0, -1, 1, -1, 1};
    for (size_t si = 0; si < peaks.size(); ++si) {
        int sr = S.seed_r[si], sc = S.seed_c[si];
        if (visited[size_t(sr) * w + sc]) continue;
        const vector<float> & G = S.grad[S.bin[size_t(sr) * w + sc]];
        float seed_val = G[size_t(sr) * w + sc];
        float min_val = (1 - kTraceTol) * seed_val;
        vector<int> px; vector<float> val;
        std::deque<int> q;
        q.push_back(sr * w + sc);
        while (!q.empty()) {
            int p = q.front(); q.pop_front();
            if (!visited[p] && G[p] > min_val) {
                px.push_back(p);
                visited[p] = 1;
                int r = p / w, c = p % w;
                for (int k = 0; k < 8; ++k) q.push_back((r + dr[k]) * w + (c + dc[k]));
            }
        }
        if (int(px.size()) > kComponentMinSize) {
            val.resize(px.size());
            for (size_t k = 0; k < px.size(); ++k) val[k] = G[px[k]];
            S.comp_px.push_back(px); S.comp_val.push_back(val);
        }
    }
    /* fit_lines_to_components line_detector.cpp:66-89 (serial order) */
    for (size_t k = 0; k < S.comp_px.size(); ++k) {
        size_t n = S.comp_px[k].size();
        vector<float> xr(n), xc(n);
        for (size_t i = 0; i < n; ++i) { xr[i] = float(S.comp_px[k][i] / w); xc[i] = float(S.comp_px[k][i] % w); }
        S.lines.push_back(fit_line_parameters(xr, xc, S.comp_val[k]));
    }
    if (!keep_planes) { for (int b = 0; b < 8; ++b) vector<float>().swap(S.grad[b]); }
}

/* ---- geometry (ref geometry.cpp) --------------------------------------------------------------------------------- */
inline float sum3(float a, float b, float c) { return a + (b + c); }   /* Eigen unrolled redux of a fixed-size-3 vector */
float seg_length(const Seg & l) { float x = l.x2 - l.x1, y = l.y2 - l.y1; return std::sqrt(x * x + y * y); }  /* :90-93 */

/* ---- f1: `refine` post-process (ref line_detector.cpp:254-444) -------------------------------------------------------
 * affinity (:357-401): for i < j, collinear if |d_i . d_j| >= 0.99 (double compare), the endpoints of the shorter
 * segment, expressed in the longer one's frame (origin = its first endpoint, axes = direction / normal, scaled by its
 * length), have |normal offset| < 0.02 (double compare) and some along-offset > -0.5 and some < 1.5.
 * labels (:274-326): graph_components follows forward edges only (u > n), so label(u) = smallest index among the nodes
 * with a forward path to u (u included).  merge (:254-271, :413-432): unique labels ascending, members in index order. */
Seg merge_lines(const vector<Seg> & g)
{
    if (g.size() == 1) return g[0];
    size_t k = g.size();
    vector<float> xr(2*k), xc(2*k), w(2*k), l(k), wts(k);
    for (size_t i = 0; i < k; ++i) {
        l[i] = seg_length(g[i]);
        wts[i] = l[i] * g[i].weight;                                             /* l.cwiseProduct(weights) */
        xr[2*i] = g[i].y1; xc[2*i] = g[i].x1; xr[2*i+1] = g[i].y2; xc[2*i+1] = g[i].x2;
        w[2*i] = wts[i]; w[2*i+1] = wts[i];
    }
    Seg m = fit_line_parameters(xr, xc, w);
    float ws = wts[0], ls = l[0];
    for (size_t i = 1; i < k; ++i) { ws = ws + wts[i]; ls = ls + l[i]; }         /* VectorXf::sum(), scalar redux */
    m.weight = ws / ls;
    return m;
}

bool refine_affine(const Seg & li, const Seg & lj, const float di[2], const float dj[2], float len_i, float len_j)
{
    float dot = di[0] * dj[0] + di[1] * dj[1];                                    /* (U^T V)(0,0) */
    if (double(std::fabs(dot)) < 0.99) return false;
    const Seg & S = (len_i < len_j) ? li : lj;                                   /* the segment whose endpoints are projected */
    const Seg & L = (len_i < len_j) ? lj : li;                                   /* the frame: first endpoint, d, n, length */
    const float * d = (len_i < len_j) ? dj : di;
    float len = (len_i < len_j) ? len_j : len_i;
    float nx = -d[1], ny = d[0];
    float ax = S.x1 - L.x1, ay = S.y1 - L.y1, bx = S.x2 - L.x1, by = S.y2 - L.y1;
    float w00 = (ax * d[0] + ay * d[1]) / len, w01 = (ax * nx + ay * ny) / len;
    float w10 = (bx * d[0] + by * d[1]) / len, w11 = (bx * nx + by * ny) / len;
    float off = std::max(std::fabs(w01), std::fabs(w11));
    if (!(double(off) < 0.02)) return false;
    return (w00 > -0.5f || w10 > -0.5f) && (w00 < 1.5f || w10 < 1.5f);
}

vector<Seg> postprocess_lines_segments(const vector<Seg> & lines)
{
    int n = int(lines.size());
    vector<float> d(2 * size_t(n)), len(n);
    for (int i = 0; i < n; ++i) {
        float x = lines[i].x2 - lines[i].x1, y = lines[i].y2 - lines[i].y1;
        len[i] = std::sqrt(x * x + y * y);
        d[2*i] = x / len[i]; d[2*i+1] = y / len[i];                               /* rowwise().normalized(): no zero check */
    }
    vector<int> label(n);
    for (int i = 0; i < n; ++i) label[i] = i;
    /* forward edges only: label(j) = min(label(j), label(i)) over edges i < j, taken in increasing i (labels of i are final) */
    for (int i = 0; i < n; ++i)
        for (int j = i + 1; j < n; ++j)
            if (refine_affine(lines[i], lines[j], &d[2*i], &d[2*j], len[i], len[j]) && label[i] < label[j]) label[j] = label[i];
    vector<Seg> res;
    for (int r = 0; r < n; ++r) {
        if (label[r] != r) continue;
        vector<Seg> g;
        for (int j = r; j < n; ++j) if (label[j] == r) g.push_back(lines[j]);
        res.push_back(merge_lines(g));
    }
    return res;
}

struct Model {           /* LinePencilModel ctor, ref line_pencil.cpp:25-32 on lines normalised as in :155-158 */
    int n;
    float bb[4], cx, cy, scale;
    vector<float> h, anchor, dir, len;   /* h: n*3; anchor, dir: n*2 */
};

void build_model(const Seg * lines, int n, Model & M)
{
    M.n = n;
    float mnx = lines[0].x1, mny = lines[0].y1, mxx = mnx, mxy = mny;      /* bounding_box geometry.cpp:96-112 */
    for (int i = 0; i < n; ++i) {
        const Seg & l = lines[i];
        mnx = std::min(mnx, std::min(l.x1, l.x2)); mxx = std::max(mxx, std::max(l.x1, l.x2));
        mny = std::min(mny, std::min(l.y1, l.y2)); mxy = std::max(mxy, std::max(l.y1, l.y2));
    }
    M.bb[0] = mnx; M.bb[1] = mny; M.bb[2] = mxx; M.bb[3] = mxy;
    float sx = mxx - mnx, sy = mxy - mny;
    M.cx = mnx + 0.5f * sx; M.cy = mny + 0.5f * sy;                        /* bbox_center :272-276 */
    M.scale = std::max(sx, sy);                                            /* bbox_size().maxCoeff() */
    M.h.resize(size_t(n) * 3); M.anchor.resize(size_t(n) * 2); M.dir.resize(size_t(n) * 2); M.len.resize(n);
    for (int i = 0; i < n; ++i) {
        float x1 = (lines[i].x1 - M.cx) / M.scale, y1 = (lines[i].y1 - M.cy) / M.scale;   /* normalize_lines :258-270 */
        float x2 = (lines[i].x2 - M.cx) / M.scale, y2 = (lines[i].y2 - M.cy) / M.scale;
        /* homogeneous :64-69 : (x1,y1,1) x (x2,y2,1) */
        float hx = y1 * 1.f - 1.f * y2, hy = 1.f * x2 - x1 * 1.f, hz = x1 * y2 - y1 * x2;
        float nrm = std::sqrt(sum3(hx * hx, hy * hy, hz * hz));
        M.h[3*i] = hx / nrm; M.h[3*i+1] = hy / nrm; M.h[3*i+2] = hz / nrm;
        M.anchor[2*i] = (x2 + x1) / 2; M.anchor[2*i+1] = (y2 + y1) / 2;
        float ddx = x2 - x1, ddy = y2 - y1;
        float ln = std::sqrt(ddx * ddx + ddy * ddy);
        M.len[i] = ln;
        M.dir[2*i] = ddx / ln; M.dir[2*i+1] = ddy / ln;
    }
}

bool sample_check(const Model & M, int a, int b)                           /* line_pencil.cpp:89-98 */
{
    float x = M.h[3*a] - M.h[3*b], y = M.h[3*a+1] - M.h[3*b+1], z = M.h[3*a+2] - M.h[3*b+2];
    return std::sqrt(sum3(x * x, y * y, z * z)) > kDegeneracyTol;
}

void fit_pair(const Model & M, int a, int b, float p[3])                   /* line_pencil.cpp:101-108 */
{
    const float * A = &M.h[3*a]; const float * B = &M.h[3*b];
    p[0] = A[1] * B[2] - A[2] * B[1];
    p[1] = A[2] * B[0] - A[0] * B[2];
    p[2] = A[0] * B[1] - A[1] * B[0];
}

/* error(): 1 - inclination, line_pencil.cpp:131-134 + geometry.cpp:214-229 */
float line_error(const Model & M, int i, const float p[3])
{
    float vx, vy;
    if (std::fabs(p[2]) < kEps) { vx = p[0]; vy = p[1]; }
    else { float px = p[0] / p[2], py = p[1] / p[2]; vx = (-M.anchor[2*i]) + px; vy = (-M.anchor[2*i+1]) + py; }
    float z = vx * vx + vy * vy;
    if (z > 0.f) { float nr = std::sqrt(z); vx = vx / nr; vy = vy / nr; }
    float inc = std::fabs(vx * M.dir[2*i] + vy * M.dir[2*i+1]);
    return -inc + 1.0f;
}

float inlier_score(const Model & M, const float p[3], float tol, const vector<int> & idx)   /* line_pencil.cpp:137-140 */
{
    float s = 0.f;
    for (size_t k = 0; k < idx.size(); ++k) {
        float t = (line_error(M, idx[k], p) < tol ? 1.f : 0.f) * M.len[idx[k]];
        s = (k == 0) ? t : s + t;
    }
    return s;
}

void fit_optimal(const Model & M, const vector<int> & idx_in, float out[3])               /* line_pencil.cpp:111-128 */
{
    vector<int> all;
    const vector<int> * idx = &idx_in;
    if (idx_in.empty()) { all.resize(M.n); for (int i = 0; i < M.n; ++i) all[i] = i; idx = &all; }
    float cov[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            float acc = 0.f;
            for (size_t k = 0; k < idx->size(); ++k) {
                int q = (*idx)[k];
                float t = (M.h[3*q+i] * M.len[q]) * M.h[3*q+j];
                acc = (k == 0) ? t : acc + t;
            }
            cov[i*3+j] = acc;
        }
    float eval[3], evec[9];
    self_adjoint_eig(cov, 3, eval, evec);
    int k = 0;
    for (int i = 1; i < 3; ++i) if (eval[i] < eval[k]) k = i;
    out[0] = evec[k*3]; out[1] = evec[k*3+1]; out[2] = evec[k*3+2];
}

/* ---- a15: mt19937 + libstdc++ generate_canonical<float,24> + choice_knuth (ref math_utils.cpp:14-39) ------------- */
struct MT19937 {
    uint32_t x[624]; int p;
    explicit MT19937(uint32_t seed) { x[0] = seed; for (int i = 1; i < 624; ++i) x[i] = 1812433253u * (x[i-1] ^ (x[i-1] >> 30)) + uint32_t(i); p = 624; }
    void twist() {
        for (int i = 0; i < 624; ++i) {
            uint32_t y = (x[i] & 0x80000000u) | (x[(i+1) % 624] & 0x7fffffffu);
            x[i] = x[(i+397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        p = 0;
    }
    uint32_t next() {
        if (p >= 624) twist();
        uint32_t y = x[p++];
        y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
        return y;
    }
};

/* k = 2^32 * generate_canonical<float,24>(rng), as an exact integer */
inline uint64_t canonical_k(uint32_t raw)
{
    uint64_t k = uint64_t(float(raw));          /* round-to-nearest-even to 24 significant bits */
    if (k >= (uint64_t(1) << 32)) k = (uint64_t(1) << 32) - 256;   /* u >= 1 -> nextafter(1,0) */
    return k;
}

void choice_knuth2(int N, MT19937 & rng, int out[2])
{
    int t = 0, m = 0;
    while (m < 2) {
        uint64_t k = canonical_k(rng.next());
        /* select unless (N-t)*u >= 2-m  <=>  (N-t)*k >= (2-m)*2^32 */
        if (uint64_t(N - t) * k >= (uint64_t(2 - m) << 32)) { t++; }
        else { out[m] = t; t++; m++; }
    }
}

/* ---- f3 (dormant in the reference): Hough-voted weights and the DirectEstimator ------------------------------------
 * libstdc++ 13 uniform_int_distribution<int>(0, m-1) over mt19937 (bits/uniform_int_dist.h: 32-bit generator range ->
 * Lemire's nearly divisionless method on a 64-bit product). */
int uniform_int_below(MT19937 & rng, uint32_t m)
{
    uint64_t product = uint64_t(rng.next()) * uint64_t(m);
    uint32_t low = uint32_t(product);
    if (low < m) {
        uint32_t threshold = uint32_t(-m) % m;
        while (low < threshold) { product = uint64_t(rng.next()) * uint64_t(m); low = uint32_t(product); }
    }
    return int(product >> 32);
}

const int kHtSpace = 65, kHtHypotheses = 20000;      /* line_pencil.h:22-23 */

/* LinePencilModel::get_weights, line_pencil.cpp:47-86; cell receives the winning accumulator cell (u, v) */
void get_weights(const Model & M, const vector<int> & idx, vector<float> & w, int cell[2])
{
    const float k = std::floor(kHtSpace / 2.f), k1 = k - 1;
    vector<float> acc(size_t(kHtSpace) * kHtSpace, 0.f);                 /* acc[u + 65 * v]: column-major like ArrayXXf */
    MT19937 rng(5489u);                                                    /* std::mt19937() default seed */
    const uint32_t m = uint32_t(idx.size());
    for (int i = 0; i < kHtHypotheses; ++i) {
        int a = idx[uniform_int_below(rng, m)];
        int b = idx[uniform_int_below(rng, m)];
        float x[3];
        fit_pair(M, a, b, x);                                              /* h.row(a).cross(h.row(b)) */
        if (std::fabs(x[0]) < 0.0001f && std::fabs(x[1]) < 0.0001f && std::fabs(x[2]) < 0.0001f) continue;
        float z = sum3(x[0] * x[0], x[1] * x[1], x[2] * x[2]);             /* normalize(): z = squaredNorm; if (z > 0) x /= sqrt(z) */
        if (z > 0.f) { float nr = std::sqrt(z); x[0] /= nr; x[1] /= nr; x[2] /= nr; }
        if (x[2] < 0.f) { x[0] = -x[0]; x[1] = -x[1]; x[2] = -x[2]; }
        int u = int(std::round(k1 * x[0] + k));
        int v = int(std::round(k1 * x[1] + k));
        acc[u + kHtSpace * v] += M.len[a] + M.len[b];
    }
    /* maxCoeff(&u, &v): Eigen's visitor walks a column-major array column by column and keeps the first strict maximum */
    int mu = 0, mv = 0; float best = acc[0];
    for (int v = 0; v < kHtSpace; ++v)
        for (int u = 0; u < kHtSpace; ++u)
            if (acc[u + kHtSpace * v] > best) { best = acc[u + kHtSpace * v]; mu = u; mv = v; }
    cell[0] = mu; cell[1] = mv;
    float p[3] = {(mu - k) / k1, (mv - k) / k1, 0.f};
    float nrm = std::sqrt(sum3(p[0] * p[0], p[1] * p[1], p[2] * p[2]));
    if (nrm > 1.f) { p[0] /= nrm; p[1] /= nrm; p[2] /= nrm; }              /* p /= p.norm(): the norm is evaluated again, same value */
    p[2] = std::sqrt(1.f - (std::pow(p[0], 2.f) + std::pow(p[1], 2.f)));
    w.resize(idx.size());
    for (size_t i = 0; i < idx.size(); ++i) {
        /* inclination(), geometry.cpp:214-229 */
        float vx, vy; const int q = idx[i];
        if (std::fabs(p[2]) < kEps) { vx = p[0]; vy = p[1]; }
        else { float px = p[0] / p[2], py = p[1] / p[2]; vx = (-M.anchor[2*q]) + px; vy = (-M.anchor[2*q+1]) + py; }
        float zz = vx * vx + vy * vy;
        if (zz > 0.f) { float nr = std::sqrt(zz); vx = vx / nr; vy = vy / nr; }
        float c = std::fabs(vx * M.dir[2*q] + vy * M.dir[2*q+1]);
        w[i] = float(std::pow(double(c), 4.0));                            /* ArrayXf::pow(int): std::pow(float, int) promotes to double */
    }
}

/* DirectEstimator::solve, estimator.h:82-96: inliers = weights > 0.95f, then fit_optimal (empty list -> all lines) */
void direct_solve(const Model & M, const vector<int> & idx, float out[3])
{
    vector<float> w; int cell[2];
    get_weights(M, idx, w, cell);
    vector<int> inl;
    for (size_t i = 0; i < idx.size(); ++i) if (w[i] > 0.95f) inl.push_back(idx[i]);
    fit_optimal(M, inl, out);
}

/* ---- f3 (dormant in the reference): PROSAC_Estimator<LinePencilModel>::solve, reference src/prosac.h:104-299 -------------
 * Progressive sampling (Chum & Matas) over the listed lines sorted by their Hough weight (argsort: std::stable_sort,
 * descending, src/utils.h:36-44).  With the class defaults (p_good_sample 0.9, max_outlier_proportion 0.5, m = 2) the
 * iteration bound is T_N = niter_RANSAC(0.9, 0.5, 2) = 9, so at most nine hypotheses are ever scored.
 * Quirks kept: the semi-random sample takes `u_n` as idx(n) — one past the paper's u_n (src/prosac.h:175) — which reads
 * beyond the list when n == N (possible only for N <= 11); this restatement requires N >= 12 like the device code. */
int niter_ransac(double p, double epsilon, int s, int Nmax)                  /* prosac.h:30-54 */
{
    if (Nmax == -1) Nmax = INT_MAX;
    if (epsilon <= 0.) return 1;
    double logarg = -std::exp(s * std::log(1. - epsilon));
    double logval = std::log(1. + logarg);
    double N = std::log(1. - p) / logval;
    if (logval < 0. && N < Nmax) return (int)std::ceil(N);
    return Nmax;
}

void choice_knuth_general(int N, int n, MT19937 & rng, int * out)            /* math_utils.cpp:14-39, any subset size */
{
    int t = 0, m = 0;
    while (m < n) {
        uint64_t k = canonical_k(rng.next());
        if (uint64_t(N - t) * k >= (uint64_t(n - m) << 32)) { t++; }
        else { out[m] = t; t++; m++; }
    }
}

struct ProsacTrace { int iterations; int n_final; int n_star; int I_N_best; int samples[9][2]; int I_N[9]; };

void prosac_solve(const Model & M, const vector<int> & indices, float tol, unsigned seed, float out[3], ProsacTrace * tr)
{
    const float eta = 0.05f, beta = 0.01f, psi = 0.02f, p_good_sample = 0.9f, max_outlier_proportion = 0.5f;   /* prosac.h:62-66 */
    static const float chi2_table[20] = { INFINITY, 6.6348966f, 5.41189443f, 4.70929225f, 4.21788459f, 3.84145882f, 3.5373846f, 3.28302029f,
        3.06490172f, 2.8743734f, 2.70554345f, 2.55422131f, 2.41732093f, 2.29250453f, 2.17795916f, 2.07225086f, 1.97422609f, 1.88294329f,
        1.79762406f, 1.71761761f };
    const float chi2_value = chi2_table[(int)std::floor(std::min(std::max(2 * psi, 0.01f), 0.2f) * 100)];          /* prosac.h:91-95,103 */
    auto Imin = [&](int m, int n) {                                             /* prosac.h:78-83 */
        double mu = n * beta;
        double sigma = std::sqrt((double)(n * beta * (1 - beta)));
        return (int)std::ceil(m + mu + sigma * std::sqrt((double)chi2_value));
    };
    vector<float> w; int cell[2];
    get_weights(M, indices, w, cell);
    const int N = int(indices.size());
    vector<int> order(N);
    for (int i = 0; i < N; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return w[a] > w[b]; });
    vector<int> idx(N);
    for (int i = 0; i < N; ++i) idx[i] = indices[order[i]];
    const int m = 2;
    const int T_N = niter_ransac(p_good_sample, max_outlier_proportion, m, -1);
    int n_star = N, I_n_star = 0, I_N_best = 0, t = 0, n = m, T_n_prime = 1, k_n_star = T_N;
    const int I_N_min = (1. - max_outlier_proportion) * N;
    double T_n = T_N;
    for (int i = 0; i < m; i++) T_n *= (double)(n - i) / (N - i);
    MT19937 rng(seed);
    float p_best[3] = {0, 0, 0};
    vector<unsigned char> best_in; bool have_best = false;
    vector<unsigned char> isIn(N);
    if (tr) { std::memset(tr, 0, sizeof *tr); }
    while (((I_N_best < I_N_min) || t <= k_n_star) && t < T_N) {
        t = t + 1;
        if ((t > T_n_prime) && (n < n_star)) {
            double T_nplus1 = (T_n * (n + 1)) / (n + 1 - m);
            n = n + 1;
            T_n_prime = T_n_prime + std::ceil(T_nplus1 - T_n);
            T_n = T_nplus1;
        }
        int sample[2];
        if (t > T_n_prime) choice_knuth_general(n, m, rng, sample);
        else { choice_knuth_general(n - 1, m - 1, rng, sample); sample[m - 1] = n; }
        if (tr) { tr->samples[t - 1][0] = sample[0]; tr->samples[t - 1][1] = sample[1]; tr->I_N[t - 1] = -1; tr->iterations = t; }
        const int a = idx[sample[0]], b = idx[sample[1]];
        if (!sample_check(M, a, b)) continue;
        float p_t[3]; fit_pair(M, a, b, p_t);
        int I_N = 0;
        for (int i = 0; i < N; ++i) { isIn[i] = line_error(M, idx[i], p_t) < tol; I_N += isIn[i]; }
        if (tr) tr->I_N[t - 1] = I_N;
        if (I_N > I_N_best) {
            I_N_best = I_N;
            p_best[0] = p_t[0]; p_best[1] = p_t[1]; p_best[2] = p_t[2];
            best_in = isIn; have_best = true;
            int n_best = N, I_n_best = I_N;
            double epsilon_n_best = (double)I_n_best / n_best;
            for (int n_test = N, I_n_test = I_N; n_test > m; n_test--) {
                if ((I_n_test * n_best > I_n_best * n_test) &&
                    (I_n_test > epsilon_n_best * n_test + std::sqrt(n_test * epsilon_n_best * (1. - epsilon_n_best) * 2.706))) {
                    if (I_n_test < Imin(m, n_test)) break;
                    n_best = n_test; I_n_best = I_n_test; epsilon_n_best = (double)I_n_best / n_best;
                }
                I_n_test -= isIn[n_test - 1];
            }
            if (I_n_best * n_star > I_n_star * n_best) {
                n_star = n_best; I_n_star = I_n_best;
                k_n_star = niter_ransac(1. - eta, 1. - I_n_star / (double)n_star, m, T_N);
            }
        }
    }
    if (tr) { tr->n_final = n; tr->n_star = n_star; tr->I_N_best = I_N_best; }
    vector<int> inl;
    if (have_best) for (int i = 0; i < N; ++i) if (best_in[i]) inl.push_back(idx[i]);
    fit_optimal(M, inl, out);                                                   /* empty list (no scored sample) -> all lines */
}

float cos_threshold(float deg) { return 1.0f - std::cos(deg / 180.f * float(M_PI)); }      /* line_pencil.cpp:143-146 */

struct Trace { int T; int * samples; float * scores; int * n_obs; int * winner; float * best_h; float * refined; };

/* RANSAC_Estimator::solve estimator.h:37-78 */
void ransac_solve(const Model & M, const vector<int> & idx, float tol, MT19937 & rng, int n_iter, float out[3], Trace * tr, int k_solve)
{
    float best_h[3] = {0.f, 0.f, 0.f};      /* uninitialised in the reference if nothing scores > 0 */
    float best = 0.f; int win = -1;
    for (int i = 0; i < n_iter; ++i) {
        int s[2];
        choice_knuth2(int(idx.size()), rng, s);
        if (tr && i < tr->T) { tr->samples[(size_t(k_solve) * tr->T + i) * 2] = s[0]; tr->samples[(size_t(k_solve) * tr->T + i) * 2 + 1] = s[1]; tr->scores[size_t(k_solve) * tr->T + i] = -1.f; }
        int a = idx[s[0]], b = idx[s[1]];
        if (!sample_check(M, a, b)) continue;
        float p[3];
        fit_pair(M, a, b, p);
        float sc = inlier_score(M, p, tol, idx);
        if (tr && i < tr->T) tr->scores[size_t(k_solve) * tr->T + i] = sc;
        if (sc > best) { best = sc; best_h[0] = p[0]; best_h[1] = p[1]; best_h[2] = p[2]; win = i; }
    }
    vector<int> inl;
    for (size_t k = 0; k < idx.size(); ++k) if (line_error(M, idx[k], best_h) < tol) inl.push_back(idx[k]);
    fit_optimal(M, inl, out);
    if (tr) { tr->winner[k_solve] = win; for (int j = 0; j < 3; ++j) { tr->best_h[k_solve*3+j] = best_h[j]; tr->refined[k_solve*3+j] = out[j]; } }
}

/* estimate_multiple_structures estimator.h:99-145 + estimate_line_pencils line_pencil.cpp:148-177 */
int estimate_line_pencils(Seg * lines, int n, Trace * tr)
{
    Model M;
    build_model(lines, n, M);
    float tol = cos_threshold(kInlierDeg), gtol = cos_threshold(kGarbageDeg);
    MT19937 rng(g_seed);
    vector<int> inlier(n, -1), garbage(n, 0);
    int num_obs = n, k = 0;
    while (num_obs >= 2 && k < kMaxModels) {
        vector<int> obs;
        for (int i = 0; i < n; ++i) if (inlier[i] < 0 && garbage[i] == 0) obs.push_back(i);
        if (tr) tr->n_obs[k] = int(obs.size());
        float hfit[3];
        ransac_solve(M, obs, tol, rng, tr ? tr->T : kRansacIter, hfit, tr, k);
        int n_in = 0, n_gb = 0;
        for (size_t j = 0; j < obs.size(); ++j) {
            float e = line_error(M, obs[j], hfit);
            if (e < tol) { inlier[obs[j]] = k; ++n_in; }
            else if (e < gtol) { garbage[obs[j]] = 1; ++n_gb; }
        }
        num_obs -= n_in + n_gb;
        ++k;
    }
    for (int i = 0; i < n; ++i) { if (garbage[i] == 1) inlier[i] = -1; lines[i].group_id = inlier[i]; }
    return k;
}

/* ---- a21: transform (ref transform.cpp, geometry.cpp:232-255, interface.cpp:93-208) ------------------------------- */
struct V3 { float x, y, z; };

V3 normalize_point(const float p[3])
{
    if (std::fabs(p[2]) < kEps) return V3{p[0], p[1], 0.f};
    return V3{p[0] / p[2], p[1] / p[2], 1.f};
}
void direction2(const V3 & a, const V3 & b, float d[2])
{
    float x = a.x - b.x * a.z, y = a.y - b.y * a.z;
    float z = x * x + y * y;
    if (z > 0.f) { float nr = std::sqrt(z); x = x / nr; y = y / nr; }
    d[0] = x; d[1] = y;
}
float distance2(const V3 & a, const V3 & b)
{
    if (a.z < kEps || b.z < kEps) return INFINITY;
    float x = a.x - b.x, y = a.y - b.y;
    return std::sqrt(x * x + y * y);
}

/* fit_vanishing_points transform.cpp:52-81; g_all: fit_single_vanishing_points :24-47 semantics when g<=0 */
std::map<int, V3> fit_vanishing_points(const Seg * lines, int n)
{
    std::map<int, V3> res;
    if (n == 0) return res;
    Model M;
    build_model(lines, n, M);
    std::set<int> groups;
    for (int i = 0; i < n; ++i) groups.insert(lines[i].group_id);
    groups.erase(-1);
    for (int g : groups) {
        vector<int> idx;
        for (int i = 0; i < n; ++i) if (lines[i].group_id == g) idx.push_back(i);
        float h[3];
        fit_optimal(M, idx, h);
        V3 vp = normalize_point(h);
        if (vp.z > 0) { vp.x = M.scale * vp.x + M.cx; vp.y = M.scale * vp.y + M.cy; }
        res[g] = vp;
    }
    return res;
}

V3 fit_single_vanishing_point(const Seg * lines, int n, int g)
{
    Model M;
    build_model(lines, n, M);
    vector<int> idx;
    if (g > 0) for (int i = 0; i < n; ++i) if (lines[i].group_id == g) idx.push_back(i);
    float h[3];
    fit_optimal(M, idx, h);
    V3 vp = normalize_point(h);
    if (vp.z > 0) { vp.x = M.scale * vp.x + M.cx; vp.y = M.scale * vp.y + M.cy; }
    return vp;
}

V3 select_vertical_point(const vector<V3> & vps, const V3 & center, float tol_deg, float min_distance)   /* :136-170 */
{
    float thr = std::cos(tol_deg / 180.0f * float(M_PI));
    for (const V3 & v : vps) {
        float d[2]; direction2(v, center, d);
        bool ang = std::fabs(d[0] * 0.f + d[1] * 1.f) > thr;
        bool dist = distance2(v, center) > min_distance;
        if (ang && dist) return v;
    }
    return V3{0, 1, 0};
}

V3 select_horizontal_point(const vector<V3> & vps, const V3 & center, const V3 & vertical, float min_distance)   /* :173-212 */
{
    float vd[2]; direction2(vertical, center, vd);
    for (const V3 & v : vps) {
        if (v.x == vertical.x && v.y == vertical.y && v.z == vertical.z) continue;
        float d[2]; direction2(v, center, d);
        float score = d[0] * vd[0] + d[1] * vd[1];
        bool horizon = score < 0.05 && score > -0.7;      /* double literals, as in the reference */
        bool dist = distance2(v, center) > min_distance;
        if (horizon && dist) return v;
    }
    return V3{1, 0, 0};
}

V3 cross3(const V3 & a, const V3 & b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

/* compute_image_transform transform.cpp:84-133; out rows: (0,0),(w,0),(w,h),(0,h) */
void compute_image_transform(int width, int height, const V3 & hvp, const V3 & vvp, float out[4][3])
{
    V3 vl = cross3(hvp, vvp);
    float H[3][3] = {{1, 0, 0}, {0, 1, 0}, {vl.x / vl.z, vl.y / vl.z, vl.z / vl.z}};
    auto mulv = [&](const V3 & v) {
        return V3{sum3(H[0][0] * v.x, H[0][1] * v.y, H[0][2] * v.z), sum3(H[1][0] * v.x, H[1][1] * v.y, H[1][2] * v.z),
                  sum3(H[2][0] * v.x, H[2][1] * v.y, H[2][2] * v.z)};
    };
    V3 hp = mulv(hvp), vp = mulv(vvp);
    if (hp.x < 0) hp = V3{-hp.x, -hp.y, -hp.z};
    if (vp.y < 0) vp = V3{-vp.x, -vp.y, -vp.z};
    float A1[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    {
        float x = hp.x, y = hp.y, z = x * x + y * y;
        if (z > 0.f) { float nr = std::sqrt(z); x = x / nr; y = y / nr; }
        A1[0][0] = x; A1[1][0] = y;
        x = vp.x; y = vp.y; z = x * x + y * y;
        if (z > 0.f) { float nr = std::sqrt(z); x = x / nr; y = y / nr; }
        A1[0][1] = x; A1[1][1] = y;
    }
    /* 3x3 inverse by cofactors (Eigen Inverse_impl.h compute_inverse_size3_helper) */
    float A[3][3];
    {
        auto cof = [&](int i, int j) {
            int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
            return A1[i1][j1] * A1[i2][j2] - A1[i1][j2] * A1[i2][j1];
        };
        float c0[3] = {cof(0, 0), cof(1, 0), cof(2, 0)};
        float det = sum3(c0[0] * A1[0][0], c0[1] * A1[1][0], c0[2] * A1[2][0]);
        float invdet = 1.f / det;
        for (int j = 0; j < 3; ++j) A[0][j] = c0[j] * invdet;
        for (int i = 1; i < 3; ++i) for (int j = 0; j < 3; ++j) A[i][j] = cof(j, i) * invdet;
    }
    float Mx[3][3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Mx[i][j] = sum3(A[i][0] * H[0][j], A[i][1] * H[1][j], A[i][2] * H[2][j]);
    float cx[4] = {0.f, float(width), float(width), 0.f}, cy[4] = {0.f, 0.f, float(height), float(height)};
    for (int k = 0; k < 4; ++k) {
        float x = cx[k] - float(width) / 2, y = cy[k] - float(height) / 2;
        float wx = sum3(Mx[0][0] * x, Mx[0][1] * y, Mx[0][2] * 1.f);
        float wy = sum3(Mx[1][0] * x, Mx[1][1] * y, Mx[1][2] * 1.f);
        float wz = sum3(Mx[2][0] * x, Mx[2][1] * y, Mx[2][2] * 1.f);
        float s = 1.0f / wz;
        out[k][0] = wx * s + float(width) / 2;
        out[k][1] = wy * s + float(height) / 2;
        out[k][2] = wz * s;
    }
}

OrcPoint pt(const V3 & v) { return OrcPoint{v.x, v.y, v.z}; }
OrcPoint pt(const float r[3]) { return OrcPoint{r[0], r[1], r[2]}; }

vector<int> g_dummy;

} // namespace

/* ================================================================================================================== */
extern "C" {

void orc_set_seed(unsigned s) { g_seed = s; }

void orc_gauss_deriv_kernel(int dir_x, float * out25) { gauss_deriv_kernel(dir_x != 0, out25); }

int orc_detect_stages(const float * img, int w, int h, int /*num_threads*/,
                      float * dx, float * dy, float * mag, int * bin, float * grad8,
                      int * seeds, int cap_seeds, int * n_seeds,
                      int * labels, int * comp_sizes, OrcLineSegment * segs, int cap_comp, int * n_comp)
{
    vector<float> im(img, img + size_t(w) * h);
    Stages S;
    detect(im, w, h, S, true);
    size_t npx = size_t(w) * h;
    if (dx)  std::memcpy(dx,  S.dx.data(), 4 * npx);
    if (dy)  std::memcpy(dy,  S.dy.data(), 4 * npx);
    if (mag) std::memcpy(mag, S.mag.data(), 4 * npx);
    if (bin) std::memcpy(bin, S.bin.data(), 4 * npx);
    if (grad8) for (int b = 0; b < 8; ++b) std::memcpy(grad8 + size_t(b) * npx, S.grad[b].data(), 4 * npx);
    *n_seeds = int(S.seed_r.size());
    if (seeds) for (size_t i = 0; i < S.seed_r.size() && int(i) < cap_seeds; ++i) { seeds[2*i] = S.seed_r[i]; seeds[2*i+1] = S.seed_c[i]; }
    *n_comp = int(S.comp_px.size());
    if (labels) {
        for (size_t i = 0; i < npx; ++i) labels[i] = -1;
        for (size_t k = 0; k < S.comp_px.size(); ++k) {
            for (int p : S.comp_px[k]) labels[p] = int(k);
            if (comp_sizes && int(k) < cap_comp) comp_sizes[k] = int(S.comp_px[k].size());
        }
    }
    if (segs) for (size_t i = 0; i < S.lines.size() && int(i) < cap_comp; ++i) segs[i] = S.lines[i];
    return 0;
}

OrcLineSegment * find_line_segment_groups(float * buffer, int width, int height, int stride, float min_length,
                                          bool refine, int /*num_threads*/, int * n_lines)
{
    vector<float> im = image_from_buffer(buffer, width, height, stride);
    Stages S;
    detect(im, width, height, S, false);
    if (S.lines.size() < 2) { *n_lines = 0; return nullptr; }                       /* interface.cpp:50-54 */
    if (refine) S.lines = postprocess_lines_segments(S.lines);                      /* interface.cpp:56-59 */
    min_length = std::max(min_length, kLineMinLength);                              /* _filter_lines interface.cpp:26-32 */
    vector<Seg> filtered;
    for (const Seg & l : S.lines) if (seg_length(l) > min_length && l.err < kLineMaxErr) filtered.push_back(l);
    if (filtered.empty()) { *n_lines = 0; return nullptr; }                         /* interface.cpp:65-69 */
    estimate_line_pencils(filtered.data(), int(filtered.size()), nullptr);
    Seg * res = new Seg[filtered.size()];
    std::copy(filtered.begin(), filtered.end(), res);
    *n_lines = int(filtered.size());
    return res;
}

void release_line_segments(OrcLineSegment ** lines) { if (*lines) { delete [] *lines; *lines = nullptr; } }

void orc_pencil_model(const OrcLineSegment * lines, int n, float * bb4, float * center2, float * scale,
                      float * h, float * anchor, float * dir, float * len)
{
    Model M; build_model(lines, n, M);
    std::memcpy(bb4, M.bb, 16); center2[0] = M.cx; center2[1] = M.cy; *scale = M.scale;
    std::memcpy(h, M.h.data(), 12 * size_t(n)); std::memcpy(anchor, M.anchor.data(), 8 * size_t(n));
    std::memcpy(dir, M.dir.data(), 8 * size_t(n)); std::memcpy(len, M.len.data(), 4 * size_t(n));
}

void orc_score_pairs(const OrcLineSegment * lines, int n, const int * idx, int n_idx, const int * pairs, int n_pairs,
                     float tol, int * check, float * hyp, float * score, float * err_out)
{
    Model M; build_model(lines, n, M);
    vector<int> ix(idx, idx + n_idx);
    for (int k = 0; k < n_pairs; ++k) {
        int a = ix[pairs[2*k]], b = ix[pairs[2*k+1]];
        check[k] = sample_check(M, a, b) ? 1 : 0;
        float p[3]; fit_pair(M, a, b, p);
        hyp[3*k] = p[0]; hyp[3*k+1] = p[1]; hyp[3*k+2] = p[2];
        score[k] = inlier_score(M, p, tol, ix);
        if (err_out && k == n_pairs - 1) for (int i = 0; i < n_idx; ++i) err_out[i] = line_error(M, ix[i], p);
    }
}

void orc_errors(const OrcLineSegment * lines, int n, const int * idx, int n_idx, const float * hyp3, float * err_out)
{
    Model M; build_model(lines, n, M);
    for (int i = 0; i < n_idx; ++i) err_out[i] = line_error(M, idx[i], hyp3);
}

void orc_fit_optimal(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float * out3)
{
    Model M; build_model(lines, n, M);
    fit_optimal(M, vector<int>(idx, idx + n_idx), out3);
}

void orc_get_weights(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float * w_out)
{
    Model M; build_model(lines, n, M);
    vector<float> w; int cell[2];
    get_weights(M, vector<int>(idx, idx + n_idx), w, cell);
    for (int i = 0; i < n_idx; ++i) w_out[i] = w[i];
}

void orc_direct_solve(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float * out3)
{
    Model M; build_model(lines, n, M);
    direct_solve(M, vector<int>(idx, idx + n_idx), out3);
}

/* PROSAC_Estimator::solve (prosac.h:104-299) on the listed lines; trace (may be null): 4 + 18 + 9 ints, see ProsacTrace */
void orc_prosac_solve(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float tol, float * out3, int * trace)
{
    Model M; build_model(lines, n, M);
    prosac_solve(M, vector<int>(idx, idx + n_idx), tol, g_seed, out3, reinterpret_cast<ProsacTrace *>(trace));
}

long long orc_choice_knuth(unsigned seed, int n, int count, int * pairs, unsigned * next_raw)
{
    MT19937 rng(seed);
    long long draws = 0;
    for (int i = 0; i < count; ++i) { choice_knuth2(n, rng, pairs + 2*i); draws += pairs[2*i+1] + 1; }
    if (next_raw) *next_raw = rng.next();
    return draws;
}

void orc_estimate_line_pencils(OrcLineSegment * lines, int n, int /*num_threads*/) { estimate_line_pencils(lines, n, nullptr); }

int orc_ransac_trace(const OrcLineSegment * lines, int n, int T, int * samples, float * scores, int * n_obs,
                     int * winner, float * best_h, float * refined, int * groups)
{
    vector<Seg> v(lines, lines + n);
    Trace tr{T, samples, scores, n_obs, winner, best_h, refined};
    for (int k = 0; k < 4; ++k) { n_obs[k] = 0; winner[k] = -1; }
    int k = estimate_line_pencils(v.data(), n, &tr);
    for (int i = 0; i < n; ++i) groups[i] = v[i].group_id;
    return k;
}

int orc_refine(const OrcLineSegment * lines, int n, OrcLineSegment * out)          /* out: capacity n; returns the merged count */
{
    vector<Seg> r = postprocess_lines_segments(vector<Seg>(lines, lines + n));
    std::copy(r.begin(), r.end(), out);
    return int(r.size());
}

void orc_fit_line_parameters(const float * X, const float * w, int n, OrcLineSegment * out)
{
    vector<float> xr(n), xc(n), ww(w, w + n);
    for (int i = 0; i < n; ++i) { xr[i] = X[2*i]; xc[i] = X[2*i+1]; }
    *out = fit_line_parameters(xr, xc, ww);
}

/* compute_rectification_transform interface.cpp:122-208 */
void orc_compute_rectification_transform(OrcLineSegment * lines, int n, int width, int height,
                                         const OrcRectificationConfig * cfg, OrcImageTransform * out)
{
    std::map<int, V3> m = fit_vanishing_points(lines, n);
    vector<V3> vps;
    for (auto & kv : m) vps.push_back(kv.second);
    V3 center{float(width) / 2, float(height) / 2, 1};
    float diag = std::sqrt(center.x * center.x + center.y * center.y);
    float min_v = std::max(cfg->vertical_vp_min_distance, 1.0f) * diag;
    V3 vp_v = select_vertical_point(vps, center, cfg->vertical_vp_angular_tolerance, min_v);
    float min_h = std::max(cfg->horizontal_vp_min_distance, 1.0f) * diag;
    V3 vp_h = select_horizontal_point(vps, center, vp_v, min_h);
    V3 v1 = vp_h; if (v1.z != 0) { v1.x -= center.x; v1.y -= center.y; }
    V3 v2 = vp_v; if (v2.z != 0) { v2.x -= center.x; v2.y -= center.y; }
    V3 v1h = v1;
    switch (cfg->h_strategy) {
        case 0: v1h.z = 0; break;                       /* ROTATE_H */
        case 1: v1h = V3{-v2.y, v2.x, 0}; break;        /* ROTATE_V */
        case 2: break;                                  /* RECTIFY */
        default: v1h = V3{1, 0, 0}; break;              /* KEEP */
    }
    V3 v2h = v2;
    switch (cfg->v_strategy) {
        case 0: v2h = V3{-v1.y, v1.x, 0}; break;
        case 1: v2h.z = 0; break;
        case 2: break;
        default: v2h = V3{0, 1, 0}; break;
    }
    float t[4][3];
    compute_image_transform(width, height, v1h, v2h, t);
    if (v1h.z != 0) { v1h.x += center.x; v1h.y += center.y; }
    if (v2h.z != 0) { v2h.x += center.x; v2h.y += center.y; }
    out->width = width; out->height = height;
    out->top_left = pt(t[0]); out->top_right = pt(t[1]); out->bottom_right = pt(t[2]); out->bottom_left = pt(t[3]);
    out->horizontal_vp = pt(v1h); out->vertical_vp = pt(v2h);
}

/* compute_rectification_transform_from_vp interface.cpp:93-119 (image centre has z=0 there) */
void orc_compute_rectification_transform_from_vp(int width, int height, const OrcPoint * vh, const OrcPoint * vv,
                                                 OrcImageTransform * out)
{
    float cx = float(width) / 2, cy = float(height) / 2;
    V3 v1{vh->x, vh->y, vh->z}; if (v1.z != 0) { v1.x -= cx; v1.y -= cy; }
    V3 v2{vv->x, vv->y, vv->z}; if (v2.z != 0) { v2.x -= cx; v2.y -= cy; }
    float t[4][3];
    compute_image_transform(width, height, v1, v2, t);
    out->width = width; out->height = height;
    out->top_left = pt(t[0]); out->top_right = pt(t[1]); out->bottom_right = pt(t[2]); out->bottom_left = pt(t[3]);
    out->horizontal_vp = *vh; out->vertical_vp = *vv;
}

void orc_fit_vanishing_point(const OrcLineSegment * lines, int n, int group, OrcPoint * out)   /* interface.cpp:211-215 */
{
    *out = pt(fit_single_vanishing_point(lines, n, group));
}

/* assign_to_group interface.cpp:218-265 */
void assign_to_group(const OrcLineSegment * lines, int n, OrcLineSegment * new_lines, int n_new, float tol_deg)
{
    std::map<int, V3> m = fit_vanishing_points(lines, n);
    vector<float> best(n_new, 0.f); vector<int> best_g(n_new, 0);
    for (auto & kv : m) {
        const V3 & p = kv.second;
        for (int i = 0; i < n_new; ++i) {
            const Seg & l = new_lines[i];
            float ax = (l.x2 + l.x1) / 2, ay = (l.y2 + l.y1) / 2;
            float dx = l.x2 - l.x1, dy = l.y2 - l.y1;
            float z = dx * dx + dy * dy;
            if (z > 0.f) { float nr = std::sqrt(z); dx = dx / nr; dy = dy / nr; }
            float vx, vy;
            if (std::fabs(p.z) < kEps) { vx = p.x; vy = p.y; }
            else { float px = p.x / p.z, py = p.y / p.z; vx = (-ax) + px; vy = (-ay) + py; }
            float zz = vx * vx + vy * vy;
            if (zz > 0.f) { float nr = std::sqrt(zz); vx = vx / nr; vy = vy / nr; }
            float inc = std::fabs(vx * dx + vy * dy);
            if (inc > best[i]) { best[i] = inc; best_g[i] = kv.first; }
        }
    }
    float thr = std::cos(tol_deg / 180 * float(M_PI));
    for (int i = 0; i < n_new; ++i) if (best[i] > thr) new_lines[i].group_id = best_g[i];
}

int orc_sizeof(int which)
{
    switch (which) { case 0: return sizeof(OrcLineSegment); case 1: return sizeof(OrcPoint); case 2: return sizeof(OrcImageTransform); case 3: return sizeof(OrcRectificationConfig); }
    return -1;
}

/* Known answers for the tie-order replay of the CUDA path (librectify_b200/csrc/k3_replay.cu): the order in which
 * std::sort with find_peaks' comparator (reference src/filter.cpp:188, element = {int i, j; float v}) leaves a value list.
 * order[j] = index into vals of the element at place j.  depth_override >= 0 calls libstdc++'s introsort loop with that
 * depth limit (then the final insertion sort), to reach the heap-sort branch on small inputs. */
void orc_std_sort_order(const float * vals, int n, int depth_override, int * order)
{
    struct Peak { int i, j; float v; };
    vector<Peak> peaks(n);
    for (int k = 0; k < n; ++k) peaks[k] = Peak{k, 0, vals[k]};
    auto cmp = [](const Peak & a, const Peak & b) { return a.v > b.v; };
    if (depth_override < 0) std::sort(peaks.begin(), peaks.end(), cmp);
    else if (n > 0) {
        std::__introsort_loop(peaks.begin(), peaks.end(), (long)depth_override, __gnu_cxx::__ops::__iter_comp_iter(cmp));
        std::__final_insertion_sort(peaks.begin(), peaks.end(), __gnu_cxx::__ops::__iter_comp_iter(cmp));
    }
    for (int k = 0; k < n; ++k) order[k] = peaks[k].i;
}
} // extern "C"
