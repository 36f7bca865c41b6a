/* librectify_b200.h — C ABI of the B200 engine underneath librectify.h.
 *
 * librectify.h is the drop-in surface (the reference's six functions).  This header is the thin C-ABI layer the
 * north star asks for between host code and the sm_100a kernels: plain pointers and sizes, no C++ or torch types.
 * It adds what the reference's one-image-per-call API cannot express on a GPU:
 *   - an explicit engine (one CUDA device, one stream, one workspace of `slots` frame slots),
 *   - batched calls over frames already resident in device memory (or in host memory),
 *   - a stage window for parity tests (every intermediate the reference computes can be fetched),
 *   - per-stage device timings for the roofline report.
 * Every function returns 0 on success and a negative code on failure; lrb_last_error() then describes it.
 * Nothing here falls back to the CPU: without a CUDA device lrb_engine_create fails.
 */
#pragma once
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
  #define LRB_API __declspec(dllexport)
#else
  #define LRB_API __attribute__((visibility("default")))
#endif

/* Same layout as librectify::LineSegment / Point / ImageTransform / RectificationConfig
 * (reference src/librectify.h:44-54,60-63,79-86,137-150); repeated as C structs so that C callers and FFI generators
 * (ctypes, cgo, JNI) can use this header without the C++ one. */
typedef struct lrb_segment { float x1, y1, x2, y2, weight, err; int group_id; } lrb_segment;
typedef struct lrb_point { float x, y, z; } lrb_point;
typedef struct lrb_transform {
    int width, height;
    lrb_point top_left, top_right, bottom_left, bottom_right, horizontal_vp, vertical_vp;
} lrb_transform;
typedef struct lrb_config {
    float vertical_vp_angular_tolerance, vertical_vp_min_distance; int v_strategy;
    float horizontal_vp_min_distance; int h_strategy;
} lrb_config;

/* Per-frame outcome of the detection + grouping pipeline. */
typedef struct lrb_frame_info {
    int n_lines;          /* filtered segments (0 when the reference would return nullptr) */
    int n_seeds;          /* NMS peaks */
    int n_components;     /* kept components = raw segments */
    int n_pixels;         /* pixels in kept components */
    int flood_rounds;     /* speculative rounds that did work */
    int flood_fallback;   /* 1 if the serial fallback flood was needed */
    int n_solves;         /* vanishing-point structures estimated (<= 4) */
    int n_obs[4];         /* observations per solve: the scoring grid is 10^4 x n_obs[k] */
    float max_magnitude;
    int floods_small;     /* tentative floods finished inside the 8-lane (256-pixel) window of k_flood_small, all rounds */
    int floods_mid;       /* ... inside the 16-lane (512-pixel) window of the second pass */
    int floods_escalated; /* tentative floods handed on to the full-tile path */
    int n_tied;           /* seeds whose magnitude equals another seed's exactly (their order comes from the std::sort replay) */
    int replay_steps;     /* partition steps replayed for that (k3_replay.cu) */
} lrb_frame_info;

typedef struct lrb_engine lrb_engine;

LRB_API const char * lrb_last_error(void);
/* Return code of the last lrb_* (or drop-in) call on the calling thread: 0 = success.  The drop-in
 * find_line_segment_groups returns nullptr both for "no segments" (reference src/interface.cpp:50-54,65-69) and for a
 * CUDA failure; this tells them apart without parsing the message. */
LRB_API int lrb_last_status(void);
LRB_API const char * lrb_version(void);

/* Seed of the sampler's mt19937 for engines created afterwards and for the drop-in API (the reference seeds from
 * std::random_device, reference src/estimator.h:35; a fixed seed — default 0 — makes the library deterministic). */
LRB_API void lrb_set_seed(unsigned seed);
/* `refine` of find_line_segment_groups (reference src/librectify.h:111-116, src/interface.cpp:56-59, src/line_detector.cpp:340-444):
 * merge collinear, overlapping raw segments before the length/err filter.  Off by default; applies to the following lrb_detect_* /
 * lrb_rectify_* calls on this engine.  The drop-in find_line_segment_groups sets it from its own `refine` argument. */
LRB_API int lrb_engine_set_refine(lrb_engine * e, int on);

/* One engine = one device + one stream + workspace for `slots` frames processed per launch group. */
LRB_API int  lrb_engine_create(int device, int slots, lrb_engine ** out);
LRB_API void lrb_engine_destroy(lrb_engine * e);
LRB_API int  lrb_engine_slots(const lrb_engine * e);

/* find_line_segment_groups for n_frames frames of equal size.
 *   frames        first pixel of frame 0; device memory for _device, host memory (pageable or pinned) for _host
 *   pitch         elements between rows, |pitch| >= width; negative: every frame pointer addresses the frame's LAST row
 *                 (bottom-up layout, reference librectify.h:15-18)
 *   frame_stride  elements between frames
 *   info          host array [n_frames]
 *   lines         host array [n_frames * lines_cap] or NULL; frame i's segments start at lines + i * lines_cap
 * A frame with more than lines_cap segments reports its true n_lines and is truncated in `lines`. */
LRB_API int lrb_detect_device(lrb_engine * e, const float * frames, int n_frames, int width, int height,
                              long long pitch, long long frame_stride, float min_length,
                              lrb_frame_info * info, lrb_segment * lines, int lines_cap);
LRB_API int lrb_detect_host(lrb_engine * e, const float * frames, int n_frames, int width, int height,
                            long long pitch, long long frame_stride, float min_length,
                            lrb_frame_info * info, lrb_segment * lines, int lines_cap);

/* Segments of slot `slot` of the engine's last launch group, copied straight from device memory: for frames whose list is
 * longer than the caller's lines_cap (replaces a second pass over the frame; reference src/interface.cpp:75-79 sizes the
 * result exactly).  Returns the number of segments copied (<= lines_cap) or a negative code. */
LRB_API int lrb_fetch_lines(lrb_engine * e, int slot, lrb_segment * lines, int lines_cap);

/* 8-bit frames: what the reference's demo does on the CPU before it calls the library —
 * cv::cvtColor(BGR2GRAY) (src/autorectify.cpp:340) and convertTo(CV_32F, 1/256.0) (src/autorectify.cpp:119) — moved to
 * the device, so a host caller ships 1 (gray) or 3 (BGR) bytes per pixel over PCIe instead of 4.  channels: 1 or 3;
 * pitch and frame_stride are in BYTES.  The result is bit-identical to passing the float frame gray/256. */
LRB_API int lrb_detect_host_u8(lrb_engine * e, const unsigned char * frames, int n_frames, int width, int height, int channels,
                               long long pitch, long long frame_stride, float min_length,
                               lrb_frame_info * info, lrb_segment * lines, int lines_cap);
LRB_API int lrb_detect_device_u8(lrb_engine * e, const unsigned char * frames, int n_frames, int width, int height, int channels,
                                 long long pitch, long long frame_stride, float min_length,
                                 lrb_frame_info * info, lrb_segment * lines, int lines_cap);
LRB_API int lrb_rectify_host_u8(lrb_engine * e, const unsigned char * frames, int n_frames, int width, int height, int channels,
                                long long pitch, long long frame_stride, float min_length,
                                const lrb_config * cfg, int n_cfg, lrb_transform * transforms, lrb_frame_info * info);
LRB_API int lrb_rectify_device_u8(lrb_engine * e, const unsigned char * frames, int n_frames, int width, int height, int channels,
                                  long long pitch, long long frame_stride, float min_length,
                                  const lrb_config * cfg, int n_cfg, lrb_transform * transforms, lrb_frame_info * info);

/* ---- the demo's image operations either side of the hot path, on the device ------------------
 * The reference delegates them to OpenCV; they are restated from OpenCV's published algorithms and pinned against cv2.
 *
 * prescale_image (reference src/autorectify.cpp:56-67): scale = min(float(max_size) / max(w, h), 1.0f) and
 * cv::resize(INTER_AREA) to cvRound(w * scale) x cvRound(h * scale).  lrb_prescale_size reports the size and scale,
 * lrb_prescale_area_f32 resizes one float frame (pitches in elements; *_on_device: the pointer is device memory). */
LRB_API int lrb_prescale_size(int width, int height, int max_size, int * dst_width, int * dst_height, float * scale);
LRB_API int lrb_prescale_area_f32(lrb_engine * e, const float * src, int src_on_device, int width, int height, long long pitch,
                                  int max_size, float * dst, int dst_on_device, long long dst_pitch);
/* homography_from_corners (reference src/autorectify.cpp:152-192): output size and the 3x3 homography (row-major
 * double[9], h33 = 1) that maps the source corners onto the transform's corners; clip = 3.0 in the demo. */
LRB_API int lrb_homography_from_corners(const lrb_transform * t, float clip, double * H9, int * dst_width, int * dst_height);
/* cv::warpPerspective(src, dst, H, Size(dst_width, dst_height)) with INTER_LINEAR and a constant 0 border (reference
 * src/autorectify.cpp:359) for 8-bit images of 1 or 3 channels; pitches in BYTES; inverse_map != 0: H9 maps
 * destination to source (cv::WARP_INVERSE_MAP). */
LRB_API int lrb_warp_perspective_u8(lrb_engine * e, const unsigned char * src, int src_on_device, int width, int height, int channels,
                                    long long pitch, const double * H9, int inverse_map, unsigned char * dst, int dst_on_device,
                                    int dst_width, int dst_height, long long dst_pitch);

/* The whole flow of the reference's demo for one 8-bit image (reference src/autorectify.cpp:113-148,330-360): gray,
 * /256, prescale to max_size (< 1: fraction of the larger side), find_line_segment_groups with min_length =
 * max(w,h)/100 of the prescaled frame, lines scaled back, compute_rectification_transform for the original size,
 * homography_from_corners(clip), warpPerspective of the input.  lines: host array [lines_cap] (truncated, *n_lines is
 * the true count); warped: host buffer of warped_bytes (needs warped_width * warped_height * channels; -5 if too small,
 * the size is reported either way).  All image work runs on the device. */
LRB_API int lrb_autorectify_u8(lrb_engine * e, const unsigned char * image, int width, int height, int channels, long long pitch,
                               float max_size, int refine, const lrb_config * cfg, float clip,
                               lrb_transform * tf, lrb_segment * lines, int lines_cap, int * n_lines,
                               unsigned char * warped, size_t warped_bytes, int * warped_width, int * warped_height);

/* detect + compute_rectification_transform per frame (the whole hot path); cfg is one config for all frames
 * (n_cfg == 1) or one per frame cycled modulo n_cfg.  transforms: host array [n_frames]. */
LRB_API int lrb_rectify_device(lrb_engine * e, const float * frames, int n_frames, int width, int height,
                               long long pitch, long long frame_stride, float min_length,
                               const lrb_config * cfg, int n_cfg, lrb_transform * transforms, lrb_frame_info * info);
LRB_API int lrb_rectify_host(lrb_engine * e, const float * frames, int n_frames, int width, int height,
                             long long pitch, long long frame_stride, float min_length,
                             const lrb_config * cfg, int n_cfg, lrb_transform * transforms, lrb_frame_info * info);

/* ---- multi-GPU batch entry (north_star: "a batch of independent images shards naturally across the 8 GPUs of
 * one box, one stream set per GPU ... at most a host-side gather of per-image ImageTransforms") ------------------------------
 * A pool owns engines_per_device engines (stream set + workspace of `slots` frames each) on every listed device.  A call
 * shards its frames: frame i of B -> device floor(i * G / B) (contiguous blocks), inside a device contiguous blocks per
 * engine, one host thread per engine (cudaSetDevice per thread), no collective; transforms and infos are written at the
 * frame's own index (the gather).  The config of frame i is cfg[i % n_cfg].  passes > 1 streams the same batch through
 * `passes` times without joining the engines in between (a benchmark's K steps); the results of the last pass are returned.
 *   _host      frames in host memory (pageable or pinned), pitch / frame_stride in elements
 *   _host_u8   8-bit gray or BGR frames in host memory, pitch / frame_stride in bytes
 *   _device    shards[d] = first pixel of the counts[d] frames resident on devices[d]; results in shard order */
typedef struct lrb_pool lrb_pool;
LRB_API int  lrb_pool_create(const int * devices, int n_devices, int engines_per_device, int slots, lrb_pool ** out);
LRB_API void lrb_pool_destroy(lrb_pool * p);
LRB_API int  lrb_pool_engines(const lrb_pool * p);
LRB_API lrb_engine * lrb_pool_engine(lrb_pool * p, int i);     /* engine i (device-major), e.g. for the stage timers */
LRB_API int lrb_pool_rectify_host(lrb_pool * p, const float * frames, int n_frames, int width, int height, long long pitch,
                                  long long frame_stride, float min_length, const lrb_config * cfg, int n_cfg,
                                  lrb_transform * transforms, lrb_frame_info * info, int passes);
LRB_API int lrb_pool_rectify_host_u8(lrb_pool * p, const unsigned char * frames, int n_frames, int width, int height, int channels,
                                     long long pitch, long long frame_stride, float min_length, const lrb_config * cfg, int n_cfg,
                                     lrb_transform * transforms, lrb_frame_info * info, int passes);
LRB_API int lrb_pool_rectify_device(lrb_pool * p, const float * const * shards, const int * counts, int width, int height,
                                    long long pitch, long long frame_stride, float min_length, const lrb_config * cfg, int n_cfg,
                                    lrb_transform * transforms, lrb_frame_info * info, int passes);

/* Host-side corner solve on caller-provided segments (pointer form of compute_rectification_transform etc.). */
LRB_API void lrb_compute_transform(const lrb_segment * lines, int n_lines, int width, int height,
                                   const lrb_config * cfg, lrb_transform * out);
LRB_API void lrb_compute_transform_from_vp(int width, int height, const lrb_point * vp_h, const lrb_point * vp_v,
                                           lrb_transform * out);
LRB_API void lrb_fit_vanishing_point(const lrb_segment * lines, int n_lines, int group, lrb_point * out);
LRB_API void lrb_assign_to_group(const lrb_segment * lines, int n_lines, lrb_segment * new_lines, int n_new_lines,
                                 float angular_tolerance_deg);

/* The two auxiliary functions on the device, for line sets large enough to be worth the trip (host arrays in,
 * host results out; same arithmetic as the host versions above, bit-identical results):
 *   lrb_fit_vanishing_point_device   reference src/interface.cpp:211-215 (group > 0: that group's lines, otherwise all lines)
 *   lrb_assign_to_group_device       reference src/interface.cpp:218-265 */
LRB_API int lrb_fit_vanishing_point_device(lrb_engine * e, const lrb_segment * lines, int n_lines, int group, lrb_point * out);
LRB_API int lrb_assign_to_group_device(lrb_engine * e, const lrb_segment * lines, int n_lines, lrb_segment * new_lines, int n_new_lines,
                                       float angular_tolerance_deg);

/* fit_vanishing_point for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same kernels and
 * bit-identical result as lrb_fit_vanishing_point_device, but
 *   - d_lines (n_lines segments) and d_out (one lrb_point) are DEVICE memory of `device`, provided by the caller;
 *   - the scratch (model of the lines: 16 bytes per line + 768) is a static __device__ array inside the library, sized for
 *     LRB_NOALLOC_MAX_LINES lines (1 MiB + 768 B, held by every GPU that loads the library until exit); more lines: error -5;
 *   - the call makes `device` current, enqueues four kernels on `stream` (a cudaStream_t of that device, NULL = default stream)
 *     and returns without synchronising: d_out is valid once the stream has passed them.  It calls no cudaMalloc / cudaFree or
 *     any allocating routine and needs no engine.
 * Every call on a device shares the one scratch array (it is re-initialised on the stream by each call): calls that overlap
 * on the GPU - different streams without an ordering between them - race; calls on one stream are safe. */
#define LRB_NOALLOC_MAX_LINES 65536
LRB_API int lrb_fit_vanishing_point_noalloc(int device, const lrb_segment * d_lines, int n_lines, int group, lrb_point * d_out,
                                            void * stream);

/* ---- the reference's dormant Hough-voted estimator ---------------------------------------------
 * LinePencilModel::get_weights (reference src/line_pencil.cpp:47-86: 20 000 pair intersections voted into a 65 x 65
 * accumulator, weight = inclination^4 towards the winning cell's direction) and DirectEstimator::solve (reference
 * src/estimator.h:82-96: fit_optimal over weights > 0.95) on the pencil model of `lines`, restricted to the observations
 * idx[0..n_idx) (NULL = all).  weights: host float[n_idx] or NULL; cell2: winning accumulator cell (u, v) or NULL;
 * vp3: the estimated vanishing point (unit vector, normalised coordinates) or NULL. */
LRB_API int lrb_direct_estimate(lrb_engine * e, const lrb_segment * lines, int n_lines, const int * idx, int n_idx,
                                float * weights, int * cell2, float * vp3);

/* PROSAC_Estimator<LinePencilModel>::solve (reference src/prosac.h:104-299; dormant in the reference like the estimator above): the
 * listed observations ordered by their Hough weight, at most nine progressively sampled hypotheses scored against all of them,
 * the reference's stopping rule, fit_optimal on the winner's inliers.  tol: inlier threshold on the line error (reference:
 * 1 - cos(2 deg)).  The generator is the engine's uniform stream (lrb_set_seed).  Needs 12 <= observations <= 46340 (below, the
 * reference reads past its list; above, its int products overflow).  vp3: unit vector in normalised coordinates;
 * trace31 (or NULL): iterations, final n, n*, best inlier count, 9 x 2 sample positions, 9 inlier counts (-1: sample rejected). */
LRB_API int lrb_prosac_estimate(lrb_engine * e, const lrb_segment * lines, int n_lines, const int * idx, int n_idx, float tol,
                                float * vp3, int * trace31);

/* ---- stage window (parity tests): contents of slot `slot` after the last lrb_detect_* call on this engine --------- */
enum lrb_stage {
    LRB_ST_DX = 0, LRB_ST_DY, LRB_ST_MAG,        /* float32 [h*w] */
    LRB_ST_BIN, LRB_ST_MASK8,                    /* uint8   [h*w] */
    LRB_ST_SEEDS,                                /* int32   [n_seeds*2] (row, col) in processing order */
    LRB_ST_LABELS,                               /* int32   [h*w] kept-component index or -1 */
    LRB_ST_COMP_SIZES,                           /* int32   [n_components] */
    LRB_ST_RAW_SEGS,                             /* lrb_segment [n_components] */
    LRB_ST_MODEL_H, LRB_ST_MODEL_ANCHOR, LRB_ST_MODEL_DIR, LRB_ST_MODEL_LEN,   /* float32 [n_lines*3|2|2|1] */
    LRB_ST_MODEL_BBOX,                           /* float32 [7]: minx miny maxx maxy cx cy scale */
    LRB_ST_PAIRS,                                /* int32   [10000*2]: samples of the LAST solve */
    LRB_ST_SCORES,                               /* float32 [4*10000] (-1 = rejected sample) */
    LRB_ST_RANSAC,                               /* int32 winner[4], float best_h[12], float refined[12] = 28 words */
    LRB_ST_STREAM,                               /* uint32 [count]: head of the uniform stream (k = 2^32 u) */
    LRB_ST_REFINED_SEGS                          /* lrb_segment [n_merged]: raw segments after the `refine` merge (engine refine on) */
};
LRB_API long long lrb_stage_fetch(lrb_engine * e, int slot, int stage, void * dst, size_t dst_bytes);

/* ---- per-stage device timings (CUDA events on the engine's stream) ---------------------------------------------- */
enum lrb_timer { LRB_T_INGEST = 0, LRB_T_GRAD, LRB_T_NMS, LRB_T_SEEDS, LRB_T_FLOOD, LRB_T_FIT, LRB_T_MODEL,
                 LRB_T_RANSAC, LRB_T_READBACK,
                 LRB_T_SCORE,      /* not a stage of its own: the k_score launches (scoring grid) inside LRB_T_RANSAC */
                 LRB_T_COUNT };
LRB_API void lrb_profile_enable(lrb_engine * e, int on);
LRB_API void lrb_profile_reset(lrb_engine * e);
/* ms[LRB_T_COUNT]: summed device milliseconds per stage; groups: launch groups timed; kernels: kernels launched */
LRB_API void lrb_profile_read(lrb_engine * e, double * ms, long long * groups, long long * kernels);
/* bytes the engine actually moved over PCIe since the last lrb_profile_reset: input frames (host entry points) and read-back */
LRB_API void lrb_profile_read_bytes(lrb_engine * e, long long * h2d_bytes, long long * d2h_bytes);
/* FP32-pipe peaks of the engine's device, measured (csrc/ubench.cu): FFMA (2 flop per instruction) and FMA-free alternating
 * FMUL / FADD (1 flop per instruction — the ceiling for this library's bit-exact, contraction-free arithmetic), in TFLOP/s.
 * The scoring kernel's roofline fraction is reported against these. */
LRB_API int lrb_ubench_fp32(lrb_engine * e, double * tflops_ffma, double * tflops_mul_add);

/* One uniform-stream sample run, for the sampler known-answer tests: `count` 2-subsets of [0,n) from the start of
 * the stream; returns the number of draws consumed (or negative on error). pairs: host int32 [count*2]. */
LRB_API long long lrb_sample_pairs(lrb_engine * e, int n, int count, int * pairs);            /* plain sequential walk */
LRB_API long long lrb_sample_pairs_parallel(lrb_engine * e, int n, int count, int * pairs);   /* speculative workers + stitch */

/* Known-answer entry point of the tie-order replay (k3_replay.cu): the reference sorts its peak list with std::sort on the
 * magnitude alone (reference src/filter.cpp:188), so the order of exact ties is what libstdc++'s introsort leaves.  `vals` is
 * a list in the order std::sort would receive it; order[j] receives the index (into vals) of the element that
 * std::sort(vals, vals + n, greater-by-value) puts at place j.  depth_override < 0: the real depth limit 2*floor(log2 n);
 * >= 0 forces the heap-sort branch early (tests).  Returns the number of partition steps replayed, negative on error. */
LRB_API long long lrb_sort_replay(lrb_engine * e, const float * vals, int n, int depth_override, int * order);

#ifdef __cplusplus
}
#endif
