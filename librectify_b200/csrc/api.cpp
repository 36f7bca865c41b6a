// api.cpp — the reference's six exported functions (reference src/interface.cpp:35-275) on top of the B200 engine.
//
// find_line_segment_groups keeps the reference's contract — synchronous, re-entrant, result array allocated with
// new[] and released with release_line_segments, nullptr + *n_lines = 0 when nothing is found — and runs the frame on
// the calling thread's current CUDA device.  Engines (stream + workspace) are pooled per device so concurrent callers
// do not serialise on one workspace.  There is no CPU fallback: any CUDA failure yields nullptr / 0 and the reason is
// available from lrb_last_error().
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/librectify.h"
#include "../../include/librectify_b200.h"

namespace {

static_assert(sizeof(librectify::LineSegment) == sizeof(lrb_segment) && sizeof(lrb_segment) == 28, "segment layout");
static_assert(sizeof(librectify::Point) == 12 && sizeof(librectify::ImageTransform) == 80 && sizeof(lrb_transform) == 80, "transform layout");
static_assert(sizeof(librectify::RectificationConfig) == 20 && sizeof(lrb_config) == 20, "config layout");

struct Pool {
    std::mutex mu;
    std::vector<std::pair<int, lrb_engine *>> idle;      // (device, engine)
    ~Pool() { for (auto & p : idle) lrb_engine_destroy(p.second); }
    lrb_engine * take(int device)
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            for (size_t i = 0; i < idle.size(); ++i)
                if (idle[i].first == device) { lrb_engine * e = idle[i].second; idle.erase(idle.begin() + i); return e; }
        }
        lrb_engine * e = nullptr;
        if (lrb_engine_create(device, 1, &e) != 0) return nullptr;
        return e;
    }
    void give(int device, lrb_engine * e) { std::lock_guard<std::mutex> lk(mu); idle.emplace_back(device, e); }
};

Pool & pool() { static Pool p; return p; }

} // namespace

namespace librectify {

extern "C" {

LineSegment * find_line_segment_groups(InputPixelType * buffer, int width, int height, int stride, float min_length,
                                       bool refine, int /*num_threads: OpenMP only*/, int * n_lines)
{
    *n_lines = 0;
    int device = 0;
    if (cudaGetDevice(&device) != cudaSuccess) return nullptr;
    lrb_engine * e = pool().take(device);
    if (!e) return nullptr;
    lrb_engine_set_refine(e, refine ? 1 : 0);                         // interface.cpp:56-59
    lrb_frame_info info;
    LineSegment * res = nullptr;
    // the first 4096 segments travel with the control block; a longer list is fetched from device memory with its exact
    // size — the frame is not processed a second time
    const int head = 4096;
    std::vector<lrb_segment> tmp(head);
    int rc = lrb_detect_host(e, buffer, 1, width, height, stride, 0, min_length, &info, tmp.data(), head);
    if (rc == 0 && info.n_lines > 0) {
        res = new LineSegment[info.n_lines];                          // interface.cpp:75
        if (info.n_lines <= head) std::memcpy(res, tmp.data(), sizeof(LineSegment) * info.n_lines);
        else if (lrb_fetch_lines(e, 0, reinterpret_cast<lrb_segment *>(res), info.n_lines) != info.n_lines) { delete[] res; res = nullptr; }
    }
    pool().give(device, e);
    if (!res) return nullptr;                                         // interface.cpp:50-54, 65-69 (or a CUDA failure: lrb_last_status())
    *n_lines = info.n_lines;
    return res;
}

void release_line_segments(LineSegment ** lines)                     // interface.cpp:268-275
{
    if (*lines != nullptr) { delete[] *lines; *lines = nullptr; }
}

ImageTransform compute_rectification_transform(LineSegment * lines, int n_lines, int width, int height, const RectificationConfig & cfg)
{
    ImageTransform T;
    lrb_compute_transform(reinterpret_cast<const lrb_segment *>(lines), n_lines, width, height,
                          reinterpret_cast<const lrb_config *>(&cfg), reinterpret_cast<lrb_transform *>(&T));
    return T;
}

ImageTransform compute_rectification_transform_from_vp(int width, int height, const Point & vp_h, const Point & vp_v)
{
    ImageTransform T;
    lrb_compute_transform_from_vp(width, height, reinterpret_cast<const lrb_point *>(&vp_h), reinterpret_cast<const lrb_point *>(&vp_v),
                                  reinterpret_cast<lrb_transform *>(&T));
    return T;
}

Point fit_vanishing_point(const LineSegment * lines, int n_lines, int group)
{
    Point p;
    lrb_fit_vanishing_point(reinterpret_cast<const lrb_segment *>(lines), n_lines, group, reinterpret_cast<lrb_point *>(&p));
    return p;
}

void assign_to_group(const LineSegment * lines, int n_lines, LineSegment * new_lines, int n_new_lines, float angular_tolerance)
{
    lrb_assign_to_group(reinterpret_cast<const lrb_segment *>(lines), n_lines, reinterpret_cast<lrb_segment *>(new_lines),
                        n_new_lines, angular_tolerance);
}

} // extern "C"

} // namespace librectify
