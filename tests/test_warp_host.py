"""Host-side parts of the demo's image operations, no GPU needed: prescale_image's size / scale rule
(reference src/autorectify.cpp:56-67) and homography_from_corners (:152-192) against cv2's golden outputs."""
import ctypes as C

import numpy as np


def _lib():
    from librectify_b200 import api
    return api, api.load_library()


def test_prescale_size_follows_cv_resize(golden):
    api, lib = _lib()
    g = golden("warp_area")
    for case in g["area_cases"]:
        k, m = str(case).split(":")
        h, w = g[f"area_src_{k}"].shape
        dw, dh, sc = C.c_int(), C.c_int(), C.c_float()
        assert lib.lrb_prescale_size(w, h, int(m), C.byref(dw), C.byref(dh), C.byref(sc)) == 0
        assert (dh.value, dw.value) == g[f"area_{k}_{m}"].shape, case
        assert np.float32(sc.value) == g[f"area_{k}_{m}_scale"], case
    assert lib.lrb_prescale_size(0, 10, 5, C.byref(dw), C.byref(dh), C.byref(sc)) != 0


def test_homography_from_corners_against_cv2(golden):
    api, lib = _lib()
    g = golden("warp_area")
    for tag in ("flow", "flow_full"):
        t = np.zeros(1, api.TF_DTYPE)
        t["width"], t["height"] = 480, 320
        c = g[f"{tag}_corners"]
        t["top_left"], t["top_right"], t["bottom_left"], t["bottom_right"] = c[0], c[1], c[2], c[3]
        H = np.zeros(9); dw, dh = C.c_int(), C.c_int()
        assert lib.lrb_homography_from_corners(t.ctypes.data, C.c_float(3.0), H.ctypes.data, C.byref(dw), C.byref(dh)) == 0
        assert (dw.value, dh.value) == tuple(int(v) for v in g[f"{tag}_dsize"])
        ref = g[f"{tag}_H"].reshape(9)
        assert np.max(np.abs(H - ref) / np.maximum(np.abs(ref), 1e-3)) < 1e-9
        # the four corner correspondences are reproduced
        Hm = H.reshape(3, 3)
        p = Hm @ np.array([480.0, 320.0, 1.0])
        xs = c[:, 0]; ys = c[:, 1]
        xmin = (xs.max() + xs.min()) / 2 - 0.5 * min(xs.max() - xs.min(), 480 * 3.0)
        assert abs(p[0] / p[2] - (c[3, 0] - xmin)) < 1e-3


def test_homography_degenerate_corners_are_rejected():
    api, lib = _lib()
    t = np.zeros(1, api.TF_DTYPE)
    t["width"], t["height"] = 640, 480                               # all four corners at the origin: no homography exists
    H = np.zeros(9); dw, dh = C.c_int(), C.c_int()
    rc = lib.lrb_homography_from_corners(t.ctypes.data, C.c_float(3.0), H.ctypes.data, C.byref(dw), C.byref(dh))
    assert rc != 0 and b"degenerate" in lib.lrb_last_error()
    assert lib.lrb_homography_from_corners(None, C.c_float(3.0), H.ctypes.data, C.byref(dw), C.byref(dh)) != 0


def test_identity_transform_gives_identity_homography():
    api, lib = _lib()
    t = np.zeros(1, api.TF_DTYPE)
    t["width"], t["height"] = 640, 480
    t["top_left"] = (0, 0, 1); t["top_right"] = (640, 0, 1); t["bottom_left"] = (0, 480, 1); t["bottom_right"] = (640, 480, 1)
    H = np.zeros(9); dw, dh = C.c_int(), C.c_int()
    assert lib.lrb_homography_from_corners(t.ctypes.data, C.c_float(3.0), H.ctypes.data, C.byref(dw), C.byref(dh)) == 0
    assert (dw.value, dh.value) == (640, 480)
    assert np.allclose(H.reshape(3, 3), np.eye(3), atol=1e-12)
