/* TEST INFRASTRUCTURE ONLY — stage-level C-ABI window onto the *compiled, unmodified* reference.
 *
 * Linked together with the reference's own nine library sources (taken where they lie under
 * /root/reference/src, never copied) into oracle/_ref/librectify_{canon,ship}.so by oracle/Makefile.
 * It only CALLS reference functions (all of them are exported C++ symbols of namespace librectify)
 * and flattens their Eigen results into plain arrays, so that tests/ and oracle/gen_golden.py can
 * pin the restatement in oracle/rectify_oracle.cpp stage by stage.
 *
 * Must be compiled with exactly the same -m / EIGEN_* flags as the reference objects it links with
 * (Eigen's allocator and vectorisation settings are part of the ABI of the Eigen types below).
 */
#include <vector>
#include <list>
#include <random>
#include <cstring>
#include <cstdint>
#include <chrono>
#include <malloc.h>
#include <Eigen/Core>
#include <Eigen/Dense>

#include "threading.h"
#include "librectify.h"
#define class struct   /* LinePencilModel keeps h/anchor/direction/length private by default; the harness reads them */
#include "line_pencil.h"
#undef class
#include "config.h"
#include "image.h"
#include "filter.h"
#include "geometry.h"
#include "line_detector.h"
#include "transform.h"
#include "math_utils.h"
#include "estimator.h"
#include <cassert>
#include <climits>
#include "utils.h"
#include "prosac.h"
#include "threading.h"
#include "utils.h"

extern "C" unsigned oracle_ref_seed_value;
unsigned oracle_ref_seed_value = 0;

namespace librectify {
/* internal reference functions that have no header declaration (reference src/line_detector.cpp:41,126,60-63,92,66) */
void image_gradients(const Image & image, Image & dx, Image & dy, Image & mag, const ThreadContext & ctx);
void gradient_directions(const Image & dx, const Image & dy, int n_bins, Image_int & grad_bin, std::vector<Image> & grad, const ThreadContext & ctx);
struct Component { Eigen::MatrixX2i px_loc; Eigen::VectorXf px_val; };
std::list<Component> find_components(std::vector<Image> & grad, const Eigen::MatrixX2i & seed, const Eigen::VectorXi & seed_bin, float tolerance);
std::vector<LineSegment> fit_lines_to_components(std::list<Component> & components, const ThreadContext & ctx);
}

using namespace librectify;
using Eigen::Index;

static Image to_image(const float * p, int w, int h)
{
    Image im(h, w);
    std::memcpy(im.data(), p, sizeof(float) * size_t(w) * size_t(h));
    return im;
}

extern "C" {

void ref_set_seed(unsigned s) { oracle_ref_seed_value = s; }

/* The reference reads uninitialised memory: grad_bin is never written where all 8 orientation energies are 0
 * (reference src/line_detector.cpp:128,152-156), and those pixels take part in the 3x3 dilation of the bin masks.
 * In a fresh process the plane comes from a fresh (zeroed) mmap, i.e. bin 0, which is what the restatement defines.
 * Pinning the mmap threshold makes every allocation >= 16 KiB a fresh zeroed mapping, so the *unmodified* reference
 * behaves like that on every call instead of only on the first one. Used for the parity oracle only, never for the
 * timed baseline.  Smaller blocks (which malloc recycles dirty) are covered by M_PERTURB = 0xff: glibc then fills
 * every malloc'ed block with the complement, i.e. zero bytes - without it the compiled reference's result was seen
 * to depend on what had been freed before the call (one segment more or less after other tests had run). */
void ref_set_deterministic_malloc() { mallopt(M_MMAP_THRESHOLD, 16384); mallopt(M_PERTURB, 0xff); }

/* 5x5 taps, row-major, dir_x != 0 -> Hx */
void ref_gauss_deriv_kernel(int dir_x, float * out25)
{
    Eigen::ArrayXXf H = gauss_deriv_kernel(EDGE_KERNEL_SIZE, EDGE_KERNEL_SIGMA, dir_x != 0);
    for (int i = 0; i < 5; ++i) for (int j = 0; j < 5; ++j) out25[i*5+j] = H(i,j);
}

/* Full detector stage dump. Any output pointer may be null.
 * seeds: cap_seeds x 2 (row,col) in the reference's sorted order; labels: h*w, index of the KEPT
 * component (list order) or -1; comp_sizes: cap_comp; segs: cap_comp LineSegment (same order). */
int ref_detect_stages(const float * img, int w, int h, int num_threads,
                      float * dx, float * dy, float * mag, int * bin, float * grad8,
                      int * seeds, int cap_seeds, int * n_seeds,
                      int * labels, int * comp_sizes, LineSegment * segs, int cap_comp, int * n_comp)
{
    ThreadContext ctx(num_threads);
    Image im = to_image(img, w, h);
    Image gx, gy, gm;
    image_gradients(im, gx, gy, gm, ctx);
    size_t npx = size_t(w) * size_t(h);
    if (dx)  std::memcpy(dx,  gx.data(), 4*npx);
    if (dy)  std::memcpy(dy,  gy.data(), 4*npx);
    if (mag) std::memcpy(mag, gm.data(), 4*npx);
    Image_int gb;
    std::vector<Image> grad;
    gradient_directions(gx, gy, 8, gb, grad, ctx);
    if (bin) std::memcpy(bin, gb.data(), 4*npx);
    if (grad8) for (int i = 0; i < 8; ++i) std::memcpy(grad8 + size_t(i)*npx, grad[i].data(), 4*npx);
    float min_seed_value = gm.maxCoeff() * (1 - std::max(std::min(float(SEED_RATIO),1.f), 0.f));
    Eigen::MatrixX2i seed;
    find_peaks(gm, SEED_DIST, min_seed_value, seed, ctx);
    *n_seeds = int(seed.rows());
    if (seeds) for (Index i = 0; i < seed.rows() && i < cap_seeds; ++i) { seeds[2*i] = seed(i,0); seeds[2*i+1] = seed(i,1); }
    Eigen::VectorXi seed_bin(seed.rows());
    for (Index i = 0; i < seed.rows(); ++i) seed_bin(i) = gb(seed(i,0), seed(i,1));
    auto comps = find_components(grad, seed, seed_bin, TRACE_TOLERANCE);
    *n_comp = int(comps.size());
    if (labels) {
        for (size_t i = 0; i < npx; ++i) labels[i] = -1;
        int k = 0;
        for (auto & c : comps) {
            for (Index j = 0; j < c.px_loc.rows(); ++j) labels[size_t(c.px_loc(j,0))*w + c.px_loc(j,1)] = k;
            if (comp_sizes && k < cap_comp) comp_sizes[k] = int(c.px_loc.rows());
            ++k;
        }
    }
    ThreadContext serial(-1);
    auto lines = fit_lines_to_components(comps, serial);
    if (segs) for (size_t i = 0; i < lines.size() && int(i) < cap_comp; ++i) segs[i] = lines[i];
    return 0;
}

/* LinePencilModel internals on the *normalised* lines exactly as estimate_line_pencils builds them
 * (reference src/line_pencil.cpp:148-166). Outputs: bb[4], center[2], scale, h[n*3], anchor[n*2], dir[n*2], len[n] */
void ref_pencil_model(const LineSegment * lines, int n, float * bb4, float * center2, float * scale,
                      float * h, float * anchor, float * dir, float * len)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    auto ln = normalize_lines(v, p, s);
    LinePencilModel model(ln);
    for (int i = 0; i < 4; ++i) bb4[i] = bb(i);
    center2[0] = p.x(); center2[1] = p.y(); *scale = s;
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < 3; ++j) h[3*i+j] = model.h(i,j);
        for (int j = 0; j < 2; ++j) { anchor[2*i+j] = model.anchor(i,j); dir[2*i+j] = model.direction(i,j); }
        len[i] = model.length(i);
    }
}

/* Per-hypothesis trace of LinePencilModel on normalised lines: for each of n_pairs index pairs (a<b, indices into
 * idx[]), reports sample_check, fit (3 floats) and inlier_score over idx[0..n_idx). err_out (optional) receives the
 * n_idx errors of the LAST pair. */
void ref_score_pairs(const LineSegment * lines, int n, const int * idx, int n_idx, const int * pairs, int n_pairs,
                     float tol, int * check, float * hyp, float * score, float * err_out)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    LinePencilModel model(normalize_lines(v, p, s));
    Eigen::ArrayXi indices = Eigen::Map<const Eigen::ArrayXi>(idx, n_idx);
    for (int k = 0; k < n_pairs; ++k) {
        Eigen::ArrayXi sample(2);
        sample << pairs[2*k], pairs[2*k+1];
        Eigen::ArrayXi si = indices(sample);
        check[k] = model.sample_check(si) ? 1 : 0;
        Eigen::Vector3f hh = model.fit(si);
        hyp[3*k] = hh.x(); hyp[3*k+1] = hh.y(); hyp[3*k+2] = hh.z();
        score[k] = model.inlier_score(hh, tol, indices);
        if (err_out && k == n_pairs - 1) {
            Eigen::ArrayXf e = model.error(hh, indices);
            for (int i = 0; i < n_idx; ++i) err_out[i] = e(i);
        }
    }
}

/* error() of an arbitrary hypothesis */
void ref_errors(const LineSegment * lines, int n, const int * idx, int n_idx, const float * hyp3, float * err_out)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    LinePencilModel model(normalize_lines(v, p, s));
    Eigen::ArrayXi indices = Eigen::Map<const Eigen::ArrayXi>(idx, n_idx);
    Eigen::Vector3f hh(hyp3[0], hyp3[1], hyp3[2]);
    Eigen::ArrayXf e = model.error(hh, indices);
    for (int i = 0; i < n_idx; ++i) err_out[i] = e(i);
}

void ref_fit_optimal(const LineSegment * lines, int n, const int * idx, int n_idx, float * out3)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    LinePencilModel model(normalize_lines(v, p, s));
    Eigen::ArrayXi indices = Eigen::Map<const Eigen::ArrayXi>(idx, n_idx);
    Eigen::Vector3f r = model.fit_optimal(indices);
    out3[0] = r.x(); out3[1] = r.y(); out3[2] = r.z();
}

/* dormant estimator: LinePencilModel::get_weights — Hough vote of 20 000 pair intersections on a
 * 65x65 accumulator, reference src/line_pencil.cpp:47-86 — and DirectEstimator::solve, reference src/estimator.h:82-96.
 * The templates are instantiated here; the reference's own build never does. */
void ref_get_weights(const LineSegment * lines, int n, const int * idx, int n_idx, float * w_out)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    LinePencilModel model(normalize_lines(v, p, s));
    Eigen::ArrayXi indices = Eigen::Map<const Eigen::ArrayXi>(idx, n_idx);
    Eigen::ArrayXf w = model.get_weights(indices);
    for (int i = 0; i < n_idx; ++i) w_out[i] = w(i);
}

void ref_direct_solve(const LineSegment * lines, int n, const int * idx, int n_idx, float * out3)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    LinePencilModel model(normalize_lines(v, p, s));
    Eigen::ArrayXi indices = Eigen::Map<const Eigen::ArrayXi>(idx, n_idx);
    DirectEstimator<LinePencilModel> est;
    Eigen::Vector3f r = est.solve(model, indices, 0.f);
    out3[0] = r.x(); out3[1] = r.y(); out3[2] = r.z();
}

/* PROSAC_Estimator<LinePencilModel>::solve (reference src/prosac.h:104-299), instantiated here; its generator is seeded through
 * the std::random_device shim (oracle/seed_shim.h), i.e. with oracle_ref_seed_value */
void ref_prosac_solve(const LineSegment * lines, int n, const int * idx, int n_idx, float tol, float * out3, int * /*trace: none*/)
{
    std::vector<LineSegment> v(lines, lines + n);
    auto bb = bounding_box(v);
    Eigen::Vector2f p = bbox_center(bb);
    float s = bbox_size(bb).maxCoeff();
    LinePencilModel model(normalize_lines(v, p, s));
    Eigen::ArrayXi indices = Eigen::Map<const Eigen::ArrayXi>(idx, n_idx);
    PROSAC_Estimator<LinePencilModel> est;
    Eigen::Vector3f r = est.solve(model, indices, tol);
    out3[0] = r.x(); out3[1] = r.y(); out3[2] = r.z();
}

/* reference src/math_utils.cpp:14 driven by std::mt19937(seed); returns total draws consumed (measured by a
 * second generator advanced in lock step is not possible, so draws = sum(s1+1)). next_raw receives rng() afterwards. */
long long ref_choice_knuth(unsigned seed, int n, int count, int * pairs, unsigned * next_raw)
{
    std::mt19937 rng(seed);
    Eigen::ArrayXi dst(2);
    long long draws = 0;
    for (int i = 0; i < count; ++i) {
        choice_knuth(n, 2, rng, dst);
        pairs[2*i] = dst(0); pairs[2*i+1] = dst(1);
        draws += dst(1) + 1;
    }
    if (next_raw) *next_raw = unsigned(rng());
    return draws;
}

void ref_estimate_line_pencils(LineSegment * lines, int n, int num_threads)
{
    std::vector<LineSegment> v(lines, lines + n);
    ThreadContext ctx(num_threads);
    estimate_line_pencils(v, MAX_MODELS, ESTIMATOR_INLIER_MAX_ANGLE_DEG, ESTIMATOR_GARBAGE_MAX_ANGLE_DEG, ctx);
    for (int i = 0; i < n; ++i) lines[i] = v[i];
}

int ref_refine(const LineSegment * lines, int n, LineSegment * out)     /* postprocess_lines_segments, serial; out: capacity n */
{
    ThreadContext ctx(-1);
    std::vector<LineSegment> r = postprocess_lines_segments(std::vector<LineSegment>(lines, lines + n), ctx);
    for (size_t i = 0; i < r.size(); ++i) out[i] = r[i];
    return int(r.size());
}

void ref_fit_line_parameters(const float * X /* n x 2 (row,col) */, const float * w, int n, LineSegment * out)
{
    Eigen::MatrixX2f Xm(n, 2);
    Eigen::VectorXf wv(n);
    for (int i = 0; i < n; ++i) { Xm(i,0) = X[2*i]; Xm(i,1) = X[2*i+1]; wv(i) = w[i]; }
    *out = fit_line_parameters(Xm, wv);
}

/* plain-pointer wrappers for the by-value / by-reference C++ signatures (ctypes cannot bind those portably) */
void ref_compute_rectification_transform(LineSegment * lines, int n, int w, int h, const RectificationConfig * cfg, ImageTransform * out)
{
    *out = compute_rectification_transform(lines, n, w, h, *cfg);
}
void ref_compute_rectification_transform_from_vp(int w, int h, const Point * vh, const Point * vv, ImageTransform * out)
{
    *out = compute_rectification_transform_from_vp(w, h, *vh, *vv);
}
void ref_fit_vanishing_point(const LineSegment * lines, int n, int group, Point * out)
{
    *out = fit_vanishing_point(lines, n, group);
}

int ref_sizeof(int which)
{
    switch (which) { case 0: return sizeof(LineSegment); case 1: return sizeof(Point); case 2: return sizeof(ImageTransform); case 3: return sizeof(RectificationConfig); }
    return -1;
}

} // extern "C"
