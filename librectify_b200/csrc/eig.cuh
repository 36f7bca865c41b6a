// eig.cuh — symmetric 2x2 / 3x3 eigen-decomposition with the conventions of the reference's linear-algebra backend.
//
// The reference obtains segment directions and vanishing points from Eigen's SelfAdjointEigenSolver
// (reference src/geometry.cpp:36-41, src/line_pencil.cpp:119-127).  The SIGN of the returned eigenvectors decides the
// endpoint order of a segment and, through it, which RANSAC samples pass sample_check — so a closed form is not
// enough: this file restates the algorithm itself (scale by max |coeff|, tridiagonalise, implicit symmetric QR with
// Wilkinson shift and Givens rotations, deflation test, ascending selection sort) in plain
// float32 with every operation rounded separately.  It is compiled for host and device so the host-side transform code
// and the kernels share one definition.
#pragma once
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define LRB_HD __host__ __device__ __forceinline__
#else
#define LRB_HD inline
#endif

namespace lrb {

struct Rot { float c, s; };

LRB_HD Rot make_givens(float p, float q)
{
    Rot g;
    if (q == 0.f) { g.c = p < 0.f ? -1.f : 1.f; g.s = 0.f; }
    else if (p == 0.f) { g.c = 0.f; g.s = q < 0.f ? 1.f : -1.f; }
    else if (fabsf(p) > fabsf(q)) {
        const float t = q / p; float u = sqrtf(1.f + t * t); if (p < 0.f) u = -u;
        g.c = 1.f / u; g.s = -t * g.c;
    } else {
        const float t = p / q; float u = sqrtf(1.f + t * t); if (q < 0.f) u = -u;
        g.s = -1.f / u; g.c = -t * g.s;
    }
    return g;
}

LRB_HD float scaled_hypot(float x, float y)
{
    x = fabsf(x); y = fabsf(y);
    const float p = fmaxf(x, y);
    if (p == 0.f) return 0.f;
    const float qp = fminf(y, x) / p;
    return p * sqrtf(1.f + qp * qp);
}

// one implicit QR step on the tridiagonal (diag, sub) block [start, end]; Q (n x n, column-major) accumulates rotations
template <int N>
LRB_HD void qr_step(float * diag, float * sub, int start, int end, float * Q)
{
    const float td = (diag[end - 1] - diag[end]) * 0.5f;
    const float e = sub[end - 1];
    float mu = diag[end];
    if (td == 0.f) mu -= fabsf(e);
    else {
        const float e2 = e * e;
        const float h = scaled_hypot(td, e);
        if (e2 == 0.f) mu -= (e / (td + (td > 0.f ? 1.f : -1.f))) * (e / h);
        else           mu -= e2 / (td + (td > 0.f ? h : -h));
    }
    float x = diag[start] - mu;
    float z = sub[start];
    for (int k = start; k < end; ++k) {
        const Rot rot = make_givens(x, z);
        const float sdk  = rot.s * diag[k] + rot.c * sub[k];
        const float dkp1 = rot.s * sub[k] + rot.c * diag[k + 1];
        diag[k]     = rot.c * (rot.c * diag[k] - rot.s * sub[k]) - rot.s * (rot.c * sub[k] - rot.s * diag[k + 1]);
        diag[k + 1] = rot.s * sdk + rot.c * dkp1;
        sub[k] = rot.c * sdk - rot.s * dkp1;
        if (k > start) sub[k - 1] = rot.c * sub[k - 1] - rot.s * z;
        x = sub[k];
        if (k < end - 1) { z = -rot.s * sub[k + 1]; sub[k + 1] = rot.c * sub[k + 1]; }
        if (!(rot.c == 1.f && rot.s == 0.f))
            for (int i = 0; i < N; ++i) {
                const float xi = Q[k * N + i], yi = Q[(k + 1) * N + i];
                Q[k * N + i]       = rot.c * xi - rot.s * yi;
                Q[(k + 1) * N + i] = rot.s * xi + rot.c * yi;
            }
    }
}

// A: symmetric N x N, row-major, only the lower triangle is read.  eval ascending; evec column-major: evec[col*N+row].
template <int N>
LRB_HD void sym_eig(const float * A, float * eval, float * evec)
{
    float m[9];
    for (int i = 0; i < 9; ++i) m[i] = 0.f;
    for (int r = 0; r < N; ++r) for (int c = 0; c <= r; ++c) m[r * N + c] = A[r * N + c];
    float scale = 0.f;
    for (int i = 0; i < N * N; ++i) scale = fmaxf(scale, fabsf(m[i]));
    if (scale == 0.f) scale = 1.f;
    for (int r = 0; r < N; ++r) for (int c = 0; c <= r; ++c) m[r * N + c] /= scale;
    float diag[3] = {0.f, 0.f, 0.f}, sub[2] = {0.f, 0.f};
    float Q[9];
    for (int i = 0; i < 9; ++i) Q[i] = 0.f;
    if (N == 2) {
        diag[0] = m[0]; diag[1] = m[3]; sub[0] = m[2];
        Q[0] = 1.f; Q[3] = 1.f;
    } else {
        diag[0] = m[0];
        const float v1norm2 = m[6] * m[6];
        if (v1norm2 <= FLT_MIN) {
            diag[1] = m[4]; diag[2] = m[8]; sub[0] = m[3]; sub[1] = m[7];
            Q[0] = Q[4] = Q[8] = 1.f;
        } else {
            const float beta = sqrtf(m[3] * m[3] + v1norm2);
            const float invBeta = 1.f / beta;
            const float m01 = m[3] * invBeta, m02 = m[6] * invBeta;
            const float q = 2.f * m01 * m[7] + m02 * (m[8] - m[4]);
            diag[1] = m[4] + m02 * q;
            diag[2] = m[8] - m02 * q;
            sub[0] = beta;
            sub[1] = m[7] - m01 * q;
            Q[0] = 1.f; Q[1 * 3 + 1] = m01; Q[2 * 3 + 1] = m02; Q[1 * 3 + 2] = m02; Q[2 * 3 + 2] = -m01;
        }
    }
    int end = N - 1, start = 0, iter = 0;
    const int maxIter = 30;
    const float precision = 2.f * FLT_EPSILON;
    while (end > 0) {
        for (int i = start; i < end; ++i)
            if (fabsf(sub[i]) <= (fabsf(diag[i]) + fabsf(diag[i + 1])) * precision || fabsf(sub[i]) <= FLT_MIN)
                sub[i] = 0.f;
        while (end > 0 && sub[end - 1] == 0.f) end--;
        if (end <= 0) break;
        iter++;
        if (iter > maxIter * N) break;
        start = end - 1;
        while (start > 0 && sub[start - 1] != 0.f) start--;
        qr_step<N>(diag, sub, start, end, Q);
    }
    if (iter <= maxIter * N) {
        for (int i = 0; i < N - 1; ++i) {
            int k = 0;
            for (int j = 1; j < N - i; ++j) if (diag[i + j] < diag[i + k]) k = j;
            if (k > 0) {
                const float t = diag[i]; diag[i] = diag[k + i]; diag[k + i] = t;
                for (int r = 0; r < N; ++r) { const float u = Q[i * N + r]; Q[i * N + r] = Q[(k + i) * N + r]; Q[(k + i) * N + r] = u; }
            }
        }
    }
    for (int i = 0; i < N; ++i) eval[i] = diag[i] * scale;
    for (int i = 0; i < N * N; ++i) evec[i] = Q[i];
}

} // namespace lrb
