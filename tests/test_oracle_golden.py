"""The CPU restatement (oracle/rectify_oracle.cpp) against golden vectors produced by the compiled reference
(tests/golden/*.npz, written by oracle/gen_golden.py).  Runs anywhere: no GPU, no /root/reference."""
import numpy as np
import pytest

from oracle.bind import RectificationConfig, transform_to_dict
from tests.util import digest, pairs_hash, tform_vec


def test_struct_sizes(oracle, golden):
    kat = golden("kat")
    assert [oracle.sizeof(i) for i in range(4)] == list(kat["sizeof"]) == [28, 12, 80, 20]


def test_taps(oracle, golden):
    kat = golden("kat")
    assert np.array_equal(oracle.gauss_deriv_kernel(True), kat["taps_x"])
    assert np.array_equal(oracle.gauss_deriv_kernel(False), kat["taps_y"])
    # SURVEY Appendix B.1
    assert float(kat["taps_x"][2, 3]) == float.fromhex("0x1.8b6582p-4")
    assert float(kat["taps_x"][0, 0]) == float.fromhex("-0x1.7e13fep-8")


def test_sampler_kats(oracle, golden):
    kat = golden("kat")
    p, _, _ = oracle.choice_knuth(0, 10, 5)
    assert np.array_equal(p, kat["knuth_n10"])
    assert p.tolist() == [[8, 9], [5, 7], [7, 9], [0, 2], [1, 6]]
    for n, want_draws in ((940, 6258234), (1570, 10418421)):
        p, d, nx = oracle.choice_knuth(0, n, 10000)
        assert np.array_equal(p[:64], kat[f"knuth_n{n}_head"])
        assert np.array_equal(p[-1], kat[f"knuth_n{n}_tail"])
        assert d == int(kat[f"knuth_n{n}_draws"]) == want_draws
        assert nx == int(kat[f"knuth_n{n}_next"])
        assert pairs_hash(p) == str(kat[f"knuth_n{n}_hash"])
    p, _, nx = oracle.choice_knuth(12345, 37, 2000)
    assert np.array_equal(p, kat["knuth_s12345_n37"]) and nx == int(kat["knuth_s12345_n37_next"])


def test_reference_test0_known_answers(oracle, golden):
    """reference src/test.cpp printed values."""
    kat = golden("kat")
    ls = kat["test0_lines"]
    vp1 = oracle.fit_vanishing_point(ls, 1)
    vp2 = oracle.fit_vanishing_point(ls, 10)
    vp3 = oracle.fit_vanishing_point(ls, -1)
    assert np.allclose(vp1, [5, 12.5, 1], atol=1e-4) and np.allclose(vp1, kat["test0_vp1"], atol=1e-5)
    assert np.allclose(vp2, [1, 0, 0], atol=1e-5) and np.allclose(vp2, kat["test0_vp2"], atol=1e-6)
    assert np.allclose(vp3, [5, 2.75789, 1], atol=1e-4) and np.allclose(vp3, kat["test0_vp3"], atol=1e-5)
    new = np.array([(5, 1, 5, 4, 1, 0, -1)], dtype=ls.dtype)
    out = oracle.assign_to_group(ls, new, 10.0)
    assert out["group_id"][0] == 1 == kat["test0_assigned"]["group_id"][0]


def test_from_vp(oracle, golden):
    kat = golden("kat")
    t = oracle.compute_rectification_transform_from_vp(1920, 1080, (3000.0, 500.0, 1.0), (900.0, -4000.0, 1.0))
    assert np.array_equal(tform_vec(transform_to_dict(t)), kat["from_vp"])
    t = oracle.compute_rectification_transform_from_vp(640, 480, (1.0, 0.0, 0.0), (0.0, 1.0, 0.0))
    assert np.array_equal(tform_vec(transform_to_dict(t)), kat["from_vp_ideal"])


def _image_of(case):
    if "image" in case:
        return case["image"]
    return case["gray_u8"].astype(np.float32) / np.float32(256.0)


@pytest.mark.parametrize("name", ["synth_320x240", "synth_640x360", "doc_image"])
def test_frame_bit_exact(oracle, golden, name):
    case = golden(name)
    img = _image_of(case)
    h, w = img.shape
    st = oracle.detect_stages(img)
    for k in ("dx", "dy", "mag", "bin", "labels", "grad8"):
        assert digest(st[k]) == str(case["sha_" + k]), k
    assert np.array_equal(st["seeds"], case["seeds"])
    assert np.array_equal(st["comp_sizes"], case["comp_sizes"])
    assert st["segs"].tobytes() == case["raw_segs"].tobytes()
    segs = oracle.find_line_segment_groups(img, float(case["min_length"]), seed=0)
    assert segs.tobytes() == case["segs"].tobytes()
    m = oracle.pencil_model(segs)
    for k in ("bb", "center", "scale", "h", "anchor", "dir", "len"):
        assert np.array_equal(m[k], case["model_" + k]), k
    hmin = 2.0 if name == "doc_image" else 1.5
    for hs in range(4):
        for vs in range(4):
            t = oracle.compute_rectification_transform(segs, w, h, RectificationConfig(40.0, 1.5, vs, hmin, hs))
            assert np.array_equal(tform_vec(transform_to_dict(t)), case["tforms"][hs * 4 + vs], equal_nan=True), (hs, vs)


def test_flipped_and_strided_input(oracle, golden):
    """reference src/librectify.h:15-18: (X, S) and (X+S*(H-1), -S) are equivalent; stride > width is honoured."""
    case = golden("synth_320x240")
    img = case["image"]
    h, w = img.shape
    pad = np.full((h, w + 7), 0.5, np.float32)
    pad[:, :w] = img
    a = oracle.find_line_segment_groups(pad, 5.0, stride=w + 7, seed=0) if False else None
    # the binding passes the array base pointer; give it the padded buffer with an explicit width
    import ctypes as C
    from oracle.bind import LineSegment, SEG_DTYPE
    lib = oracle.lib
    for flipped in (False, True):
        oracle.set_seed(0)
        n = C.c_int(0)
        stride = w + 7
        base = pad.ctypes.data + (4 * stride * (h - 1) if flipped else 0)
        ptr = lib.find_line_segment_groups(C.cast(C.c_void_p(base), C.POINTER(C.c_float)), w, h,
                                           -stride if flipped else stride, C.c_float(5.0), False, -1, C.byref(n))
        out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_byte)), shape=(n.value * 28,)).view(SEG_DTYPE).copy()
        pp = C.cast(ptr, C.POINTER(LineSegment)); lib.release_line_segments(C.byref(pp))
        assert out.tobytes() == case["segs"].tobytes()


def test_degenerate_inputs(oracle):
    flat = np.full((64, 96), 0.5, np.float32)
    assert oracle.find_line_segment_groups(flat, 5.0) is None            # reference interface.cpp:50-54
    tiny = np.zeros((5, 5), np.float32); tiny[2, 2] = 1
    assert oracle.find_line_segment_groups(tiny, 5.0) is None
    step = np.zeros((40, 40), np.float32); step[:, 20:] = 1.0
    out = oracle.find_line_segment_groups(step, 100.0)                    # all filtered out: interface.cpp:65-69
    assert out is None
