"""`refine` post-process (reference src/line_detector.cpp:254-444, src/interface.cpp:56-59) on the GPU,
through the C ABI, against the oracle: merged raw segments and the final grouped segments are compared BIT-EXACT."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from librectify_b200 import api as a
    a.load_library()
    return a


def _synth(seed, w, h):
    from librectify_b200.synth import manhattan_frame
    return manhattan_frame(seed, w, h).numpy()


def _cases(golden):
    out = []
    for name in ("synth_320x240", "synth_640x360", "doc_image"):
        g = golden(name)
        img = g["image"] if "image" in g else g["gray_u8"].astype(np.float32) / np.float32(256.0)
        out.append((name, img, float(g["min_length"])))
    out.append(("synth5_1280x720", _synth(5, 1280, 720), 12.8))
    out.append(("synth1_1920x1080", _synth(1, 1920, 1080), 19.2))
    return out


def test_refined_segments_and_groups_match_oracle(api, oracle, golden):
    eng = api.Engine(0, 1)
    eng.set_refine(True)
    for name, img, ml in _cases(golden):
        info, lines = eng.detect(img, ml)
        raw = eng.fetch("raw_segs")
        merged = eng.fetch("refined_segs")
        ref_merged = oracle.refine(raw)
        assert len(merged) == len(ref_merged) < len(raw), name              # every fixture has collinear pieces to merge
        assert merged.tobytes() == ref_merged.tobytes(), name
        ref = oracle.find_line_segment_groups(img, ml, refine=True, seed=0)
        assert ref is not None and lines[0].tobytes() == ref.tobytes(), name
    # the flag is per engine and can be switched off again
    eng.set_refine(False)
    name, img, ml = _cases(golden)[1]
    _, lines = eng.detect(img, ml)
    assert lines[0].tobytes() == oracle.find_line_segment_groups(img, ml, refine=False, seed=0).tobytes()
    eng.close()


def test_refine_against_reference_golden(api, golden):
    """the committed fixture was produced by the compiled reference itself (oracle/gen_golden.py)"""
    g = golden("refine")
    eng = api.Engine(0, 1)
    eng.set_refine(True)
    _, lines = eng.detect(g["image"], float(g["min_length"]))
    assert eng.fetch("refined_segs").tobytes() == g["merged"].tobytes()
    assert lines[0].tobytes() == g["lines_refined"].tobytes()
    eng.close()


def test_drop_in_refine_flag(api, oracle, golden):
    g = golden("synth_640x360")
    img = g["image"]; ml = float(g["min_length"])
    a = api.find_line_segment_groups(img, ml, refine=True)
    b = api.find_line_segment_groups(img, ml, refine=False)
    assert a.tobytes() == oracle.find_line_segment_groups(img, ml, refine=True, seed=0).tobytes()
    assert b.tobytes() == oracle.find_line_segment_groups(img, ml, refine=False, seed=0).tobytes()
    assert len(a) != len(b)


def test_refine_in_a_batch(api, oracle):
    frames = np.stack([_synth(s, 640, 360) for s in (3, 4, 5)])
    eng = api.Engine(0, 3)
    eng.set_refine(True)
    info, lines = eng.detect(frames, 6.4)
    for i in range(3):
        assert lines[i].tobytes() == oracle.find_line_segment_groups(frames[i], 6.4, refine=True, seed=0).tobytes(), i
    eng.close()
