/* TEST INFRASTRUCTURE ONLY — C ABI of the CPU restatement of librectify's hot path.
 *
 * This is the parity oracle of the repository: a plain C++ (no Eigen, no OpenMP) restatement of
 * find_line_segment_groups + compute_rectification_transform of RomanJuranek/librectify, each function citing the
 * reference file:line it follows.  It is pinned against the compiled reference itself (oracle/_ref, built from the
 * unmodified sources by oracle/Makefile) in tests/test_oracle_vs_ref.py and against the committed fixtures in
 * tests/golden/.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it; the product
 * library (librectify_b200/csrc) never includes, links or calls anything in this directory.
 */
#pragma once
#include <cstdint>

extern "C" {

struct OrcLineSegment { float x1, y1, x2, y2, weight, err; int group_id; };   /* reference src/librectify.h:44-54 */
struct OrcPoint { float x, y, z; };                                            /* reference src/librectify.h:60-63 */
struct OrcImageTransform {                                                     /* reference src/librectify.h:79-86 */
    int width, height;
    OrcPoint top_left, top_right, bottom_left, bottom_right, horizontal_vp, vertical_vp;
};
struct OrcRectificationConfig {                                                /* reference src/librectify.h:137-150 */
    float vertical_vp_angular_tolerance, vertical_vp_min_distance; int v_strategy;
    float horizontal_vp_min_distance; int h_strategy;
};

/* public API, same names/ABI as the reference (src/librectify.h:111-222); the by-value/by-reference C++ signatures
 * are additionally exposed as orc_* pointer wrappers, mirroring oracle/ref_harness.cpp */
OrcLineSegment * find_line_segment_groups(float * buffer, int width, int height, int stride, float min_length,
                                          bool refine, int num_threads, int * n_lines);
void release_line_segments(OrcLineSegment ** lines);
void assign_to_group(const OrcLineSegment * lines, int n, OrcLineSegment * new_lines, int n_new, float tol_deg);

void orc_set_seed(unsigned s);
void orc_gauss_deriv_kernel(int dir_x, float * out25);
int  orc_detect_stages(const float * img, int w, int h, int num_threads,
                       float * dx, float * dy, float * mag, int * bin, float * grad8,
                       int * seeds, int cap_seeds, int * n_seeds,
                       int * labels, int * comp_sizes, OrcLineSegment * segs, int cap_comp, int * n_comp);
void orc_pencil_model(const OrcLineSegment * lines, int n, float * bb4, float * center2, float * scale,
                      float * h, float * anchor, float * dir, float * len);
void orc_score_pairs(const OrcLineSegment * lines, int n, const int * idx, int n_idx, const int * pairs, int n_pairs,
                     float tol, int * check, float * hyp, float * score, float * err_out);
void orc_errors(const OrcLineSegment * lines, int n, const int * idx, int n_idx, const float * hyp3, float * err_out);
void orc_fit_optimal(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float * out3);
/* dormant estimator: LinePencilModel::get_weights (line_pencil.cpp:47-86), DirectEstimator::solve (estimator.h:82-96) */
void orc_get_weights(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float * w_out);
void orc_direct_solve(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float * out3);
/* dormant estimator: PROSAC_Estimator<LinePencilModel>::solve (prosac.h:104-299); trace: 31 ints or null */
void orc_prosac_solve(const OrcLineSegment * lines, int n, const int * idx, int n_idx, float tol, float * out3, int * trace);
long long orc_choice_knuth(unsigned seed, int n, int count, int * pairs, unsigned * next_raw);
void orc_estimate_line_pencils(OrcLineSegment * lines, int n, int num_threads);
void orc_fit_line_parameters(const float * X, const float * w, int n, OrcLineSegment * out);
int  orc_refine(const OrcLineSegment * lines, int n, OrcLineSegment * out);   /* postprocess_lines_segments, line_detector.cpp:340-444 */
void orc_compute_rectification_transform(OrcLineSegment * lines, int n, int w, int h,
                                         const OrcRectificationConfig * cfg, OrcImageTransform * out);
void orc_compute_rectification_transform_from_vp(int w, int h, const OrcPoint * vh, const OrcPoint * vv,
                                                 OrcImageTransform * out);
void orc_fit_vanishing_point(const OrcLineSegment * lines, int n, int group, OrcPoint * out);
int  orc_sizeof(int which);
/* order in which std::sort (find_peaks' comparator, reference src/filter.cpp:188) leaves `vals`: known answers for k3_replay.cu */
void orc_std_sort_order(const float * vals, int n, int depth_override, int * order);

/* extra trace used by the GPU parity tests: full RANSAC trace of one estimate_line_pencils call.
 * samples: [4][T][2] positions into the observation list of each solve, scores: [4][T] (-1 = rejected by
 * sample_check), n_obs[4], winner[4] (-1 = none), best_h[4][3], refined[4][3]; returns number of solves run. */
int orc_ransac_trace(const OrcLineSegment * lines, int n, int T, int * samples, float * scores, int * n_obs,
                     int * winner, float * best_h, float * refined, int * groups);
}
