// transform_host.cpp — the part of the hot path that stays on the host: vanishing-point refit per group, selection of
// the vertical / horizontal point, the 15 rectify / rotate / keep strategies and the corner solve.
//
// Mirrors reference src/transform.cpp:24-212, src/interface.cpp:93-265 and the helpers in src/geometry.cpp:232-255.
// It is ~0.2 ms of scalar float32 work on <= 4 points per frame, so the north star keeps it on the
// CPU; it is compiled with -ffp-contract=off so the float32 results are those of the reference's operation order.
#include <algorithm>
#include <cmath>
#include <map>
#include <set>
#include <vector>

#include "../../include/librectify_b200.h"
#include "eig.cuh"

namespace lrb {

namespace {

const float kEpsH = 1e-6f;     // reference src/config.h:10

struct Vec3 { float x, y, z; };

inline float add3(float a, float b, float c) { return a + (b + c); }   // reduction order of a fixed-size 3-vector

// Pencil model of a set of segments in bounding-box-normalised coordinates (src/line_pencil.cpp:25-32, 148-158)
struct Pencil {
    int n = 0;
    float cx = 0, cy = 0, scale = 1;
    std::vector<float> h, len;   // h: 3 per line

    Pencil(const lrb_segment * s, int count) : n(count), h(size_t(count) * 3), len(count)
    {
        float x0 = s[0].x1, y0 = s[0].y1, x1 = x0, y1 = y0;
        for (int i = 0; i < n; ++i) {
            x0 = std::min(x0, std::min(s[i].x1, s[i].x2)); x1 = std::max(x1, std::max(s[i].x1, s[i].x2));
            y0 = std::min(y0, std::min(s[i].y1, s[i].y2)); y1 = std::max(y1, std::max(s[i].y1, s[i].y2));
        }
        const float sx = x1 - x0, sy = y1 - y0;
        cx = x0 + 0.5f * sx; cy = y0 + 0.5f * sy; scale = std::max(sx, sy);
        for (int i = 0; i < n; ++i) {
            const float ax = (s[i].x1 - cx) / scale, ay = (s[i].y1 - cy) / scale;
            const float bx = (s[i].x2 - cx) / scale, by = (s[i].y2 - cy) / scale;
            const float hx = ay * 1.f - 1.f * by, hy = 1.f * bx - ax * 1.f, hz = ax * by - ay * bx;
            const float nrm = std::sqrt(add3(hx * hx, hy * hy, hz * hz));
            h[3 * i] = hx / nrm; h[3 * i + 1] = hy / nrm; h[3 * i + 2] = hz / nrm;
            const float dx = bx - ax, dy = by - ay;
            len[i] = std::sqrt(dx * dx + dy * dy);
        }
    }

    // least-squares common point of the indexed lines; an empty index list means all lines (src/line_pencil.cpp:111-128)
    Vec3 common_point(const std::vector<int> & which) const
    {
        float cov[9];
        const int cnt = which.empty() ? n : int(which.size());
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                float acc = 0.f;
                for (int k = 0; k < cnt; ++k) {
                    const int q = which.empty() ? k : which[k];
                    const float t = (h[3 * q + r] * len[q]) * h[3 * q + c];
                    acc = k ? acc + t : t;
                }
                cov[r * 3 + c] = acc;
            }
        float eval[3], evec[9];
        sym_eig<3>(cov, eval, evec);
        int k = 0;
        for (int i = 1; i < 3; ++i) if (eval[i] < eval[k]) k = i;
        return Vec3{evec[3 * k], evec[3 * k + 1], evec[3 * k + 2]};
    }

    // homogeneous -> image coordinates (src/geometry.cpp:232-238, src/transform.cpp:41-45)
    Vec3 to_image(const Vec3 & p) const
    {
        Vec3 q = std::fabs(p.z) < kEpsH ? Vec3{p.x, p.y, 0.f} : Vec3{p.x / p.z, p.y / p.z, 1.f};
        if (q.z > 0) { q.x = scale * q.x + cx; q.y = scale * q.y + cy; }
        return q;
    }
};

void unit_direction(const Vec3 & a, const Vec3 & b, float d[2])           // src/geometry.cpp:241-246
{
    float x = a.x - b.x * a.z, y = a.y - b.y * a.z;
    const float z = x * x + y * y;
    if (z > 0.f) { const float nr = std::sqrt(z); x = x / nr; y = y / nr; }
    d[0] = x; d[1] = y;
}

float point_distance(const Vec3 & a, const Vec3 & b)                      // src/geometry.cpp:249-255
{
    if (a.z < kEpsH || b.z < kEpsH) return INFINITY;
    const float x = a.x - b.x, y = a.y - b.y;
    return std::sqrt(x * x + y * y);
}

std::map<int, Vec3> group_points(const lrb_segment * s, int n)            // src/transform.cpp:52-81
{
    std::map<int, Vec3> out;
    if (n <= 0) return out;
    const Pencil P(s, n);
    std::set<int> ids;
    for (int i = 0; i < n; ++i) if (s[i].group_id != -1) ids.insert(s[i].group_id);
    for (int g : ids) {
        std::vector<int> which;
        for (int i = 0; i < n; ++i) if (s[i].group_id == g) which.push_back(i);
        out[g] = P.to_image(P.common_point(which));
    }
    return out;
}

Vec3 pick_vertical(const std::vector<Vec3> & vps, const Vec3 & c, float tol_deg, float min_dist)   // src/transform.cpp:136-170
{
    const float thr = std::cos(tol_deg / 180.0f * float(M_PI));
    for (const Vec3 & v : vps) {
        float d[2]; unit_direction(v, c, d);
        if (std::fabs(d[0] * 0.f + d[1] * 1.f) > thr && point_distance(v, c) > min_dist) return v;
    }
    return Vec3{0, 1, 0};
}

Vec3 pick_horizontal(const std::vector<Vec3> & vps, const Vec3 & c, const Vec3 & vert, float min_dist)  // src/transform.cpp:173-212
{
    float vd[2]; unit_direction(vert, c, vd);
    for (const Vec3 & v : vps) {
        if (v.x == vert.x && v.y == vert.y && v.z == vert.z) continue;
        float d[2]; unit_direction(v, c, d);
        const float s = d[0] * vd[0] + d[1] * vd[1];
        if (s < 0.05 && s > -0.7 && point_distance(v, c) > min_dist) return v;      // double literals as in the reference
    }
    return Vec3{1, 0, 0};
}

inline Vec3 cross(const Vec3 & a, const Vec3 & b) { return Vec3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// src/transform.cpp:84-133: H sends the vanishing line to infinity, A aligns the two directions with the axes, the
// four corners (0,0) (w,0) (w,h) (0,h) are mapped about the image centre.
void corner_solve(int w, int h, const Vec3 & hv, const Vec3 & vv, float out[4][3])
{
    const Vec3 vl = cross(hv, vv);
    const float H[3][3] = {{1, 0, 0}, {0, 1, 0}, {vl.x / vl.z, vl.y / vl.z, vl.z / vl.z}};
    auto apply = [&](const Vec3 & v) {
        return Vec3{add3(H[0][0] * v.x, H[0][1] * v.y, H[0][2] * v.z), add3(H[1][0] * v.x, H[1][1] * v.y, H[1][2] * v.z),
                    add3(H[2][0] * v.x, H[2][1] * v.y, H[2][2] * v.z)};
    };
    Vec3 hp = apply(hv), vp = apply(vv);
    if (hp.x < 0) hp = Vec3{-hp.x, -hp.y, -hp.z};
    if (vp.y < 0) vp = Vec3{-vp.x, -vp.y, -vp.z};
    float B[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    auto unit2 = [](float & x, float & y) { const float z = x * x + y * y; if (z > 0.f) { const float nr = std::sqrt(z); x = x / nr; y = y / nr; } };
    { float x = hp.x, y = hp.y; unit2(x, y); B[0][0] = x; B[1][0] = y; }
    { float x = vp.x, y = vp.y; unit2(x, y); B[0][1] = x; B[1][1] = y; }
    // inverse by cofactors, first column scaled via the determinant (Eigen's fixed-size 3x3 inverse)
    float A[3][3];
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return B[i1][j1] * B[i2][j2] - B[i1][j2] * B[i2][j1];
    };
    const float c0[3] = {cof(0, 0), cof(1, 0), cof(2, 0)};
    const float inv = 1.f / add3(c0[0] * B[0][0], c0[1] * B[1][0], c0[2] * B[2][0]);
    for (int j = 0; j < 3; ++j) A[0][j] = c0[j] * inv;
    for (int i = 1; i < 3; ++i) for (int j = 0; j < 3; ++j) A[i][j] = cof(j, i) * inv;
    float M[3][3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) M[i][j] = add3(A[i][0] * H[0][j], A[i][1] * H[1][j], A[i][2] * H[2][j]);
    const float X[4] = {0.f, float(w), float(w), 0.f}, Y[4] = {0.f, 0.f, float(h), float(h)};
    for (int k = 0; k < 4; ++k) {
        const float x = X[k] - float(w) / 2, y = Y[k] - float(h) / 2;
        const float px = add3(M[0][0] * x, M[0][1] * y, M[0][2] * 1.f);
        const float py = add3(M[1][0] * x, M[1][1] * y, M[1][2] * 1.f);
        const float pz = add3(M[2][0] * x, M[2][1] * y, M[2][2] * 1.f);
        const float s = 1.0f / pz;
        out[k][0] = px * s + float(w) / 2; out[k][1] = py * s + float(h) / 2; out[k][2] = pz * s;
    }
}

inline lrb_point P3(const float r[3]) { return lrb_point{r[0], r[1], r[2]}; }
inline lrb_point P3(const Vec3 & v) { return lrb_point{v.x, v.y, v.z}; }

void pack_corners(lrb_transform * T, int w, int h, const float t[4][3])
{
    T->width = w; T->height = h;
    T->top_left = P3(t[0]); T->top_right = P3(t[1]); T->bottom_right = P3(t[2]); T->bottom_left = P3(t[3]);
}

} // namespace

} // namespace lrb

using namespace lrb;

extern "C" {

// src/interface.cpp:122-208
void lrb_compute_transform(const lrb_segment * lines, int n, int w, int h, const lrb_config * cfg, lrb_transform * out)
{
    std::vector<Vec3> vps;
    for (auto & kv : group_points(lines, n)) vps.push_back(kv.second);            // ascending group id
    const Vec3 c{float(w) / 2, float(h) / 2, 1};
    const float diag = std::sqrt(c.x * c.x + c.y * c.y);
    const Vec3 vv = pick_vertical(vps, c, cfg->vertical_vp_angular_tolerance, std::max(cfg->vertical_vp_min_distance, 1.0f) * diag);
    const Vec3 hv = pick_horizontal(vps, c, vv, std::max(cfg->horizontal_vp_min_distance, 1.0f) * diag);
    Vec3 v1 = hv; if (v1.z != 0) { v1.x -= c.x; v1.y -= c.y; }
    Vec3 v2 = vv; if (v2.z != 0) { v2.x -= c.x; v2.y -= c.y; }
    Vec3 a = v1, b = v2;
    switch (cfg->h_strategy) {                    // 0 ROTATE_H, 1 ROTATE_V, 2 RECTIFY, 3 KEEP (src/librectify.h:126-132)
        case 0: a.z = 0; break;
        case 1: a = Vec3{-v2.y, v2.x, 0}; break;
        case 2: break;
        default: a = Vec3{1, 0, 0}; break;
    }
    switch (cfg->v_strategy) {
        case 0: b = Vec3{-v1.y, v1.x, 0}; break;
        case 1: b.z = 0; break;
        case 2: break;
        default: b = Vec3{0, 1, 0}; break;
    }
    float t[4][3];
    corner_solve(w, h, a, b, t);
    if (a.z != 0) { a.x += c.x; a.y += c.y; }
    if (b.z != 0) { b.x += c.x; b.y += c.y; }
    pack_corners(out, w, h, t);
    out->horizontal_vp = P3(a); out->vertical_vp = P3(b);
}

// src/interface.cpp:93-119
void lrb_compute_transform_from_vp(int w, int h, const lrb_point * ph, const lrb_point * pv, lrb_transform * out)
{
    const float cx = float(w) / 2, cy = float(h) / 2;
    Vec3 a{ph->x, ph->y, ph->z}; if (a.z != 0) { a.x -= cx; a.y -= cy; }
    Vec3 b{pv->x, pv->y, pv->z}; if (b.z != 0) { b.x -= cx; b.y -= cy; }
    float t[4][3];
    corner_solve(w, h, a, b, t);
    pack_corners(out, w, h, t);
    out->horizontal_vp = *ph; out->vertical_vp = *pv;
}

// src/interface.cpp:211-215 -> src/transform.cpp:24-47 (group <= 0 selects every line, as the reference's `g > 0` does)
void lrb_fit_vanishing_point(const lrb_segment * lines, int n, int group, lrb_point * out)
{
    const Pencil P(lines, n);
    std::vector<int> which;
    if (group > 0) for (int i = 0; i < n; ++i) if (lines[i].group_id == group) which.push_back(i);
    *out = P3(P.to_image(P.common_point(which)));
}

// src/interface.cpp:218-265
void lrb_assign_to_group(const lrb_segment * lines, int n, lrb_segment * fresh, int n_fresh, float tol_deg)
{
    const std::map<int, Vec3> vps = group_points(lines, n);
    std::vector<float> best(n_fresh, 0.f); std::vector<int> best_g(n_fresh, 0);
    for (auto & kv : vps) {
        const Vec3 & p = kv.second;
        for (int i = 0; i < n_fresh; ++i) {
            const lrb_segment & l = fresh[i];
            const float ax = (l.x2 + l.x1) / 2, ay = (l.y2 + l.y1) / 2;
            float dx = l.x2 - l.x1, dy = l.y2 - l.y1;
            { const float z = dx * dx + dy * dy; if (z > 0.f) { const float nr = std::sqrt(z); dx = dx / nr; dy = dy / nr; } }
            float vx, vy;
            if (std::fabs(p.z) < kEpsH) { vx = p.x; vy = p.y; }
            else { const float px = p.x / p.z, py = p.y / p.z; vx = (-ax) + px; vy = (-ay) + py; }
            { const float z = vx * vx + vy * vy; if (z > 0.f) { const float nr = std::sqrt(z); vx = vx / nr; vy = vy / nr; } }
            const float inc = std::fabs(vx * dx + vy * dy);
            if (inc > best[i]) { best[i] = inc; best_g[i] = kv.first; }
        }
    }
    const float thr = std::cos(tol_deg / 180 * float(M_PI));
    for (int i = 0; i < n_fresh; ++i) if (best[i] > thr) fresh[i].group_id = best_g[i];
}

} // extern "C"
