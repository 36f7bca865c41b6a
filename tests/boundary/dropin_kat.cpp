// Boundary check of the drop-in header: a plain C++ program that includes include/librectify.h exactly as a user of the
// reference includes src/librectify.h, links against librectify_b200.so and exercises the host-side functions of the six
// exported symbols.  The inputs are the hand-made quadrilateral of the reference's own smoke program (reference
// src/test.cpp:18-39); what that program prints there are the known answers checked by the test.
// (find_line_segment_groups is referenced so that the link also proves that symbol; it needs a GPU and is not called.)
#include <cstdio>

#include "librectify.h"

using namespace librectify;

static void show(const char * name, const Point & p) { std::printf("%s=[%g,%g,%g]\n", name, p.x, p.y, p.z); }

int main(int argc, char **)
{
    LineSegment quad[4] = {
        {0.f, 0.f, 10.f, 0.f, 1.f, 0.f, 10},      // bottom edge, group 10
        {10.f, 0.f, 8.f, 5.f, 1.f, 0.f, 1},       // right edge,  group 1
        {8.f, 5.f, 2.f, 5.f, 1.f, 0.f, 10},       // top edge,    group 10
        {2.f, 5.f, 0.f, 0.f, 1.f, 0.f, 1},        // left edge,   group 1
    };
    show("vp1", fit_vanishing_point(quad, 4, 1));
    show("vp2", fit_vanishing_point(quad, 4, 10));
    show("vp3", fit_vanishing_point(quad, 4, -1));
    LineSegment probe = {5.f, 1.f, 5.f, 4.f, 1.f, 0.f, -1};
    assign_to_group(quad, 4, &probe, 1, 10.f);
    std::printf("Assigned to group %d\n", probe.group_id);
    RectificationConfig cfg;                        // the header's default member initialisers
    ImageTransform t = compute_rectification_transform(quad, 4, 10, 5, cfg);
    std::printf("size=%dx%d strategies=%d,%d\n", t.width, t.height, (int)cfg.h_strategy, (int)cfg.v_strategy);
    ImageTransform u = compute_rectification_transform_from_vp(640, 480, Point{1.f, 0.f, 0.f}, Point{0.f, 1.f, 0.f});
    std::printf("ideal=[%g,%g] [%g,%g]\n", u.top_left.x, u.top_left.y, u.bottom_right.x, u.bottom_right.y);
    if (argc > 1000) {                              // never taken: only makes the linker resolve the GPU entry points
        int n = 0; LineSegment * r = find_line_segment_groups(nullptr, 0, 0, 0, 0.f, false, -1, &n); release_line_segments(&r);
    }
    return 0;
}
