"""CPU-side checks of the product library: it loads, exports every symbol the headers declare, keeps the reference's
struct layouts, refuses to run without a CUDA device (no CPU fallback), and its host-only functions (the corner solve
that the north star keeps on the CPU) reproduce the golden vectors.  No GPU compute happens here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from librectify_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(api.LIB_PATH):
        from librectify_b200 import build
        build.build()
    return api.load_library()


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = "\n".join(l for l in txt.splitlines() if not l.lstrip().startswith("#"))
    names = re.findall(r"\b(?:LRB_API|DLL_PUBLIC)\b[^;{]*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", txt)
    return sorted(set(names))


def test_exports_every_declared_symbol(lib):
    ref_api = _declared("librectify.h")
    assert ref_api == sorted(["find_line_segment_groups", "release_line_segments", "compute_rectification_transform",
                              "compute_rectification_transform_from_vp", "fit_vanishing_point", "assign_to_group"])
    ext = _declared("librectify_b200.h")
    assert len(ext) >= 19 and all(n.startswith("lrb_") for n in ext)
    for name in ref_api + ext:
        assert hasattr(lib, name), name


def test_struct_layouts(golden):
    kat = golden("kat")
    assert [api.SEG_DTYPE.itemsize, 12, api.TF_DTYPE.itemsize, api.CFG_DTYPE.itemsize] == list(kat["sizeof"])
    assert (api.ROTATE_H, api.ROTATE_V, api.RECTIFY, api.KEEP) == (0, 1, 2, 3)      # reference src/librectify.h:126-132
    c = api.make_config()[0]
    assert (c["vertical_vp_angular_tolerance"], c["vertical_vp_min_distance"], c["v_strategy"],
            c["horizontal_vp_min_distance"], c["h_strategy"]) == (40.0, 1.5, 2, 1.5, 2)


def test_no_cpu_fallback(lib):
    """Without a CUDA device the engine must refuse to exist and the drop-in call must return nullptr / 0."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    assert lib.lrb_engine_create(0, 1, C.byref(h)) != 0 and not h.value
    assert b"no CPU path" in lib.lrb_last_error()
    with pytest.raises(api.LibrectifyError):
        api.Engine(0, 1)
    img = np.zeros((32, 32), np.float32)
    n = C.c_int(5)
    assert not lib.find_line_segment_groups(img.ctypes.data, 32, 32, 32, C.c_float(5.0), False, -1, C.byref(n))
    assert n.value == 0


def test_product_does_not_touch_the_oracle():
    """The shipped sources never include, link or import anything under oracle/."""
    pkg = os.path.join(ROOT, "librectify_b200")
    for dirpath, _, files in os.walk(pkg):
        if "_obj" in dirpath or "__pycache__" in dirpath:
            continue
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".inc", ".py")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "rectify_oracle" not in txt and "oracle.bind" not in txt and "from oracle" not in txt, f


# ---- host-side functions (they run on the CPU in the product too) -------------------------------------------------------
def test_test0_known_answers(golden):
    """reference src/test.cpp printed values."""
    kat = golden("kat")
    ls = kat["test0_lines"]
    assert np.allclose(api.fit_vanishing_point(ls, 1), [5, 12.5, 1], atol=1e-4)
    assert np.allclose(api.fit_vanishing_point(ls, 10), [1, 0, 0], atol=1e-5)
    assert np.allclose(api.fit_vanishing_point(ls, -1), [5, 2.75789, 1], atol=1e-4)
    for g, key in ((1, "test0_vp1"), (10, "test0_vp2"), (-1, "test0_vp3")):
        assert np.array_equal(np.float32(api.fit_vanishing_point(ls, g)), kat[key])
    new = np.array([(5, 1, 5, 4, 1, 0, -1)], dtype=ls.dtype)
    out = api.assign_to_group(ls, new, 10.0)
    assert out["group_id"][0] == 1 == kat["test0_assigned"]["group_id"][0]


def test_from_vp(golden):
    kat = golden("kat")
    t = api.compute_rectification_transform_from_vp(1920, 1080, (3000.0, 500.0, 1.0), (900.0, -4000.0, 1.0))
    assert np.array_equal(api.tform_vec(t), kat["from_vp"])
    t = api.compute_rectification_transform_from_vp(640, 480, (1.0, 0.0, 0.0), (0.0, 1.0, 0.0))
    assert np.array_equal(api.tform_vec(t), kat["from_vp_ideal"])


@pytest.mark.parametrize("name", ["synth_320x240", "synth_640x360", "doc_image"])
def test_corner_solve_all_strategies(golden, name):
    """compute_rectification_transform on the golden segments: all 16 h/v strategy pairs, bit-exact."""
    case = golden(name)
    h, w = (case["image"] if "image" in case else case["gray_u8"]).shape
    hmin = 2.0 if name == "doc_image" else 1.5
    for hs in range(4):
        for vs in range(4):
            t = api.compute_rectification_transform(case["segs"], w, h, api.make_config(40.0, 1.5, vs, hmin, hs))
            assert np.array_equal(api.tform_vec(t), case["tforms"][hs * 4 + vs], equal_nan=True), (hs, vs)


def test_corner_solve_matches_oracle_on_perturbed_groups(oracle, golden):
    """random regroupings / subsets of the golden segments: product host code == oracle restatement"""
    from oracle.bind import RectificationConfig, transform_to_dict
    from tests.util import tform_vec
    case = golden("synth_640x360")
    segs = case["segs"].copy()
    h, w = case["image"].shape
    rng = np.random.default_rng(3)
    for trial in range(12):
        s = segs[rng.random(len(segs)) < 0.6].copy()
        if trial % 3 == 0:
            s["group_id"] = rng.integers(-1, 4, len(s))
        if trial == 5:
            s["group_id"] = -1
        cfg = (float(rng.uniform(10, 60)), float(rng.uniform(0.5, 3)), int(rng.integers(0, 4)), float(rng.uniform(0.5, 3)), int(rng.integers(0, 4)))
        a = api.tform_vec(api.compute_rectification_transform(s, w, h, api.make_config(*cfg)))
        b = tform_vec(transform_to_dict(oracle.compute_rectification_transform(s, w, h, RectificationConfig(*cfg))))
        assert np.array_equal(a, b, equal_nan=True), trial
        for g in (-1, 0, 1, 2):
            assert np.array_equal(np.float32(api.fit_vanishing_point(s, g)), np.float32(oracle.fit_vanishing_point(s, g)), equal_nan=True)


def test_cpp_program_links_against_the_drop_in(tmp_path):
    """A C++ translation unit that includes include/librectify.h the way a user of the reference includes src/librectify.h
    (reference src/test.cpp:1-43 does exactly this) compiles with g++, links against librectify_b200.so and prints the
    reference program's four known answers.  Host-side functions only: no GPU needed."""
    import shutil
    import subprocess
    cxx = shutil.which("g++")
    assert cxx, "g++ is part of the image"
    src = os.path.join(ROOT, "tests", "boundary", "dropin_kat.cpp")
    exe = str(tmp_path / "dropin_kat")
    libdir = os.path.dirname(api.LIB_PATH)
    subprocess.check_call([cxx, "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), src, "-o", exe,
                           "-L", libdir, "-l:librectify_b200.so", "-Wl,-rpath," + libdir])
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.splitlines()
    assert out[0] == "vp1=[5,12.5,1]"
    assert out[1] == "vp2=[1,0,0]"
    assert out[2].startswith("vp3=[5,2.7578") and out[2].endswith(",1]")
    assert out[3] == "Assigned to group 1"
    assert out[4] == "size=10x5 strategies=2,2"                      # RECTIFY / RECTIFY defaults (reference src/librectify.h:137-150)
    assert out[5] == "ideal=[0,0] [640,480]"                         # ideal points on both axes: the identity
