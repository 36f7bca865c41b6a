"""The oracle's restatement of the `refine` post-process (reference src/line_detector.cpp:254-444) against the fixture the
compiled reference produced (tests/golden/refine.npz, oracle/gen_golden.py) and, where oracle/_ref is built, against the
compiled reference live.  Bit-exact."""
import numpy as np


def test_oracle_refine_matches_reference_golden(oracle, golden):
    g = golden("refine")
    assert oracle.refine(g["raw"]).tobytes() == g["merged"].tobytes()
    ml = float(g["min_length"])
    assert oracle.find_line_segment_groups(g["image"], ml, refine=True, seed=0).tobytes() == g["lines_refined"].tobytes()
    assert oracle.find_line_segment_groups(g["image"], ml, refine=False, seed=0).tobytes() == g["lines_plain"].tobytes()
    assert len(g["merged"]) < len(g["raw"]) and len(g["lines_refined"]) != len(g["lines_plain"])


def test_oracle_refine_matches_compiled_reference(oracle, ref_canon, golden):
    from librectify_b200.synth import manhattan_frame
    for seed, w, h in ((3, 640, 360), (9, 517, 389)):
        img = manhattan_frame(seed, w, h).numpy()
        raw = oracle.detect_stages(img, want_grad8=False)["segs"]
        assert oracle.refine(raw).tobytes() == ref_canon.refine(raw).tobytes()
        a = oracle.find_line_segment_groups(img, 6.0, refine=True, seed=0)
        b = ref_canon.find_line_segment_groups(img, 6.0, refine=True, num_threads=-1, seed=0)
        assert a.tobytes() == b.tobytes()


def test_refine_edge_cases(oracle):
    from oracle.bind import SEG_DTYPE
    # a single segment and two unrelated segments pass through unchanged
    one = np.array([(0, 0, 10, 0, 1, 0.1, -1)], SEG_DTYPE)
    assert oracle.refine(one).tobytes() == one.tobytes()
    two = np.array([(0, 0, 10, 0, 1, 0.1, -1), (0, 5, 0, 25, 2, 0.2, -1)], SEG_DTYPE)
    assert oracle.refine(two).tobytes() == two.tobytes()
    # forward-edge labelling: 0-2 and 1-2 are collinear pairs, 0-1 is not (gap too long for the shorter frame) ->
    # segment 2 joins the group of 0, segment 1 stays alone (graph_components only follows u > n, line_detector.cpp:293)
    chain = np.array([(0, 0, 10, 0, 1, 0, -1), (40, 0, 50, 0, 1, 0, -1), (12, 0, 38, 0, 1, 0, -1)], SEG_DTYPE)
    out = oracle.refine(chain)
    assert len(out) == 2
    assert out[1].tobytes() == chain[1].tobytes()
    assert min(out[0]["x1"], out[0]["x2"]) == 0 and max(out[0]["x1"], out[0]["x2"]) == 38
