"""The demo's image operations either side of the hot path, on the device (reference
src/autorectify.cpp:56-67 prescale_image, :152-192 homography_from_corners, :355-360 warpPerspective), through the C ABI.
The reference delegates them to OpenCV; tests/golden/warp_area.npz holds cv2 4.13's outputs (oracle/gen_golden_cv.py) and,
for the whole flow, the compiled reference's lines and corners on cv2's prescaled frame.
Bars: warpPerspective and INTER_AREA bit-exact against cv2 for the same matrix / frame; the homography within 1e-9
relative of cv2.findHomography (cv2 solves the same four-point system by a normalised DLT); the whole flow: lines and
corners bit-exact against the reference, warped image equal to cv2's except where a source coordinate lands within
1e-9 px of a 1/32-px rounding boundary (bounded at 1e-4 of the pixels, observed 0)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from librectify_b200 import api as a
    a.load_library()
    return a


def test_warp_perspective_matches_cv2_bit_exact(api, golden):
    g = golden("warp_area")
    eng = api.Engine(0, 1)
    for i in range(3):
        H = g[f"warp{i}_H"]; ds = tuple(int(v) for v in g[f"warp{i}_dsize"])
        assert np.array_equal(eng.warp_perspective(g["warp_bgr"], H, ds), g[f"warp{i}_bgr"]), i
        assert np.array_equal(eng.warp_perspective(g["warp_gray"], H, ds), g[f"warp{i}_gray"]), i
        assert np.array_equal(eng.warp_perspective(g["warp_gray"], H, ds, inverse_map=True), g[f"warp{i}_gray_inv"]), i
    eng.close()


def test_warp_device_pointers_and_pitches(api, golden):
    import torch
    g = golden("warp_area")
    eng = api.Engine(0, 1)
    H = np.ascontiguousarray(g["warp1_H"], np.float64); dw, dh = (int(v) for v in g["warp1_dsize"])
    src = torch.zeros((240, 1024), dtype=torch.uint8, device="cuda"); src[:, :960] = torch.from_numpy(g["warp_bgr"].reshape(240, 960)).cuda()
    dst = torch.full((dh, 2048), 7, dtype=torch.uint8, device="cuda")
    rc = eng.lib.lrb_warp_perspective_u8(eng.h, src.data_ptr(), 1, 320, 240, 3, 1024, H.ctypes.data, 0, dst.data_ptr(), 1, dw, dh, 2048)
    assert rc == 0, eng.lib.lrb_last_error()
    torch.cuda.synchronize()
    out = dst.cpu().numpy()
    assert np.array_equal(out[:, : dw * 3].reshape(dh, dw, 3), g["warp1_bgr"])
    assert (out[:, dw * 3:] == 7).all()                              # nothing written beyond a row
    bad = eng.lib.lrb_warp_perspective_u8(eng.h, src.data_ptr(), 1, 320, 240, 2, 1024, H.ctypes.data, 0, dst.data_ptr(), 1, dw, dh, 2048)
    assert bad != 0
    Z = np.zeros(9); rc = eng.lib.lrb_warp_perspective_u8(eng.h, src.data_ptr(), 1, 320, 240, 3, 1024, Z.ctypes.data, 0, dst.data_ptr(), 1, dw, dh, 2048)
    assert rc != 0 and b"singular" in eng.lib.lrb_last_error()
    eng.close()


def test_prescale_area_matches_cv2_bit_exact(api, golden):
    g = golden("warp_area")
    eng = api.Engine(0, 1)
    for case in g["area_cases"]:
        k, m = str(case).split(":")
        src = g[f"area_src_{k}"].astype(np.float32) / np.float32(256.0)
        out, scale = eng.prescale(src, int(m))
        ref = g[f"area_{k}_{m}"]
        assert scale == g[f"area_{k}_{m}_scale"], case
        assert out.shape == ref.shape and out.tobytes() == ref.tobytes(), case
    out, _ = eng.prescale(g["area_float_src"], 100)                  # arbitrary floats, fractional scale: order of operations matters
    assert out.tobytes() == g["area_float_100"].tobytes()
    eng.close()


def test_homography_from_corners_against_cv2(api, golden):
    g = golden("warp_area")
    eng = api.Engine(0, 1)
    for tag in ("flow", "flow_full"):
        t = np.zeros(1, api.TF_DTYPE)
        t["width"], t["height"] = 480, 320
        c = g[f"{tag}_corners"]
        t["top_left"], t["top_right"], t["bottom_left"], t["bottom_right"] = c[0], c[1], c[2], c[3]
        H, ds = eng.homography_from_corners(t, 3.0)
        assert ds == tuple(int(v) for v in g[f"{tag}_dsize"])
        assert np.max(np.abs(H - g[f"{tag}_H"]) / np.maximum(np.abs(g[f"{tag}_H"]), 1e-3)) < 1e-9
    eng.close()


@pytest.mark.parametrize("tag", ["flow", "flow_full"])
def test_whole_demo_flow(api, golden, tag):
    g = golden("warp_area")
    eng = api.Engine(0, 1)
    warped, tf, lines = eng.autorectify(g["flow_src"], max_size=float(g[f"{tag}_max_size"]))
    assert lines.tobytes() == g[f"{tag}_lines"].tobytes()
    c = np.array([tf["top_left"], tf["top_right"], tf["bottom_left"], tf["bottom_right"]], np.float32)
    assert c.tobytes() == g[f"{tag}_corners"].tobytes()
    ref = g[f"{tag}_warped"]
    assert warped.shape == ref.shape
    diff = warped != ref
    assert diff.mean() <= 1e-4, diff.mean()
    assert np.abs(warped.astype(int) - ref.astype(int)).max() <= 8
    eng.close()
