/* librectify.h — drop-in public interface of the B200-native librectify hot path.
 *
 * This header declares exactly the six entry points and five plain-data types that the reference library
 * exports (reference src/librectify.h:44-222, implemented in reference src/interface.cpp:35-275), with identical
 * names, argument order, struct layouts (28 / 12 / 80 / 20 bytes) and enum values, so an application built
 * against the reference header links against librectify_b200's shared object unchanged.
 *
 * What differs is only where the work runs: find_line_segment_groups() uploads the frame to the current B200 and
 * runs hand-written sm_100a kernels (librectify_b200/csrc); there is no CPU fallback — without a usable CUDA
 * device the call fails (nullptr, *n_lines = 0, reason in lrb_last_error(), see librectify_b200.h).
 *
 * Image layout (reference src/librectify.h:8-18): a grayscale float32 raster given by a pointer, width, height and
 * a row stride counted in ELEMENTS.  The stride may be negative, in which case `buffer` addresses the row that is
 * highest in memory: (X, S) and (X + S*(H-1), -S) describe the same image.
 */
#pragma once

#if defined(_WIN32) || defined(_WINDOWS)
  #ifdef LIBRECTIFY_EXPORTS
    #define DLL_PUBLIC __declspec(dllexport)
  #else
    #define DLL_PUBLIC __declspec(dllimport)
  #endif
#else
  #define DLL_PUBLIC __attribute__((visibility("default")))
#endif

namespace librectify {

/* One detected segment (reference src/librectify.h:44-54).  28 bytes, no padding. */
struct LineSegment
{
    float x1, y1, x2, y2;   /* endpoints in pixel coordinates (x = column, y = row) */
    float weight;           /* mean orientation energy of the supporting pixels */
    float err;              /* mean absolute normal residual of the supporting pixels */
    int   group_id;         /* pencil (vanishing point) index 0..3, or -1 when unassigned */
};

/* Homogeneous 2-D point (reference src/librectify.h:60-63).  z == 0 denotes a direction. */
struct Point
{
    float x, y, z;
};

/* Where the four image corners move under the rectifying homography, in a frame whose centre stays fixed
 * (reference src/librectify.h:79-86).  80 bytes. */
struct ImageTransform
{
    int   width;
    int   height;
    Point top_left, top_right, bottom_left, bottom_right;
    Point horizontal_vp;
    Point vertical_vp;
};

using InputPixelType = float;

/* Detect segments and group them into at most four pencils (reference src/librectify.h:111-116,
 * src/interface.cpp:35-80).
 *   min_length  segments not longer than max(min_length, 5) px, or with err >= 2, are dropped
 *   refine      merge collinear, overlapping raw segments before the filter (src/line_detector.cpp:340-444), as the reference does
 *   num_threads accepted and ignored (the reference's OpenMP thread count; the GPU path has no use for it)
 *   n_lines     receives the element count of the returned array
 * Returns an array owned by the caller (free it with release_line_segments), or nullptr with *n_lines = 0
 * when fewer than two raw segments were found or none survived the filter. */
extern "C" DLL_PUBLIC LineSegment * find_line_segment_groups(
    InputPixelType * buffer, int width, int height, int stride,
    float min_length, bool refine, int num_threads, int * n_lines);

/* delete[] the array and null the pointer (reference src/librectify.h:123, src/interface.cpp:268-275). */
extern "C" DLL_PUBLIC void release_line_segments(LineSegment ** lines);

/* Per-axis behaviour of compute_rectification_transform (reference src/librectify.h:126-132). */
enum RectificationStrategy
{
    ROTATE_H,   /* rotate so that the horizontal vanishing direction becomes horizontal */
    ROTATE_V,   /* rotate so that the vertical vanishing direction becomes vertical */
    RECTIFY,    /* full perspective correction along this axis */
    KEEP,       /* leave this axis alone */
};

/* Reference src/librectify.h:137-150.  20 bytes. */
struct RectificationConfig
{
    float vertical_vp_angular_tolerance {40};        /* degrees around the image's vertical axis */
    float vertical_vp_min_distance {1.5f};           /* in units of the image half-diagonal */
    RectificationStrategy v_strategy {RECTIFY};
    float horizontal_vp_min_distance {1.5f};         /* in units of the image half-diagonal */
    RectificationStrategy h_strategy {RECTIFY};
};

/* Pick a vertical and a horizontal vanishing point among the groups and return the warped image corners
 * (reference src/librectify.h:175-178, src/interface.cpp:122-208).  Runs on the host. */
extern "C" DLL_PUBLIC ImageTransform compute_rectification_transform(
    LineSegment * lines, int n_lines, int width, int height, const RectificationConfig & cfg);

/* Same corner solve from two given vanishing points (reference src/librectify.h:181-183,
 * src/interface.cpp:93-119). */
extern "C" DLL_PUBLIC ImageTransform compute_rectification_transform_from_vp(
    int width, int height, const Point & vp_h, const Point & vp_v);

/* Least-squares vanishing point of one group; group <= 0 fits all lines (reference src/librectify.h:199-201,
 * src/interface.cpp:211-215, src/transform.cpp:24-47 — note the `g > 0` test there). */
extern "C" DLL_PUBLIC Point fit_vanishing_point(const LineSegment * lines, int n_lines, int group);

/* Give each of `new_lines` the group of the best-aligned vanishing point of `lines` if it is within
 * angular_tolerance degrees (reference src/librectify.h:219-222, src/interface.cpp:218-265). */
extern "C" DLL_PUBLIC void assign_to_group(
    const LineSegment * lines, int n_lines, LineSegment * new_lines, int n_new_lines, float angular_tolerance);

} // namespace librectify
