"""TEST INFRASTRUCTURE ONLY — writes tests/golden/warp_area.npz: the image operations the reference's demo delegates to
OpenCV (prescale_image, findHomography, warpPerspective; reference src/autorectify.cpp:56-67,
152-192,355-360), produced by cv2 (OpenCV 4.13, this container) and, for the whole-flow case, by the COMPILED, UNMODIFIED
reference (oracle/_ref/librectify_canon.so) on cv2's prescaled frame.

    python -m oracle.gen_golden_cv

OpenCV is a third-party dependency of the demo and is not under /root/reference; these vectors are what pins the CUDA
restatement of its algorithms (librectify_b200/csrc/k_warp.cu) on the GPU box.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def prescale_cv(img, max_size):
    """prescale_image, src/autorectify.cpp:56-67 (float scale exactly as the C++ evaluates it)"""
    import cv2
    w = max(img.shape[1], img.shape[0])
    scale = min(np.float32(max_size) / np.float32(w), np.float32(1.0))
    if scale == 1.0:
        return img, scale
    return cv2.resize(img, None, fx=float(scale), fy=float(scale), interpolation=cv2.INTER_AREA), scale


def homography_from_corners_cv(t, clip):
    """src/autorectify.cpp:152-192 with cv2.findHomography; t: dict with width, height and the four corners (x, y, z)"""
    import cv2
    f = np.float32
    x = [f(t[k][0]) for k in ("top_left", "top_right", "bottom_left", "bottom_right")]
    y = [f(t[k][1]) for k in ("top_left", "top_right", "bottom_left", "bottom_right")]
    xmin, xmax, ymin, ymax = min(x), max(x), min(y), max(y)
    xc = f((xmax + xmin) / f(2)); yc = f((ymax + ymin) / f(2))
    width = min(f(xmax - xmin), f(f(t["width"]) * f(clip))); height = min(f(ymax - ymin), f(f(t["height"]) * f(clip)))
    dsize = (int(width), int(height))
    xmin = f(xc - f(f(0.5) * width)); ymin = f(yc - f(f(0.5) * height))
    src = np.array([[0, 0], [t["width"], 0], [0, t["height"]], [t["width"], t["height"]]], np.float32)
    dst = np.array([[f(x[i] - xmin), f(y[i] - ymin)] for i in range(4)], np.float32)
    H, _ = cv2.findHomography(src, dst)
    return H, dsize


def main():
    import cv2
    from oracle.bind import Ref, RectificationConfig, transform_to_dict
    rng = np.random.default_rng(20261104)
    out = {"cv2_version": np.array(cv2.__version__)}
    # ---- warpPerspective: BGR and gray, forward and inverse maps, sources partly outside ---------------------------
    bgr = cv2.imread("/root/reference/doc/image.jpg")[60:300, 150:470].copy()            # 240 x 320 x 3
    gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
    Hs = [np.array([[1.05, 0.08, -12.3], [0.03, 0.97, 7.7], [1.2e-4, -0.8e-4, 1.0]]),
          np.array([[0.7, -0.25, 60.0], [0.2, 0.8, -30.0], [-3e-4, 2e-4, 1.0]]),
          np.array([[1.0, 0.0, 0.5], [0.0, 1.0, -0.25], [0.0, 0.0, 1.0]])]
    dsizes = [(300, 250), (411, 203), (320, 240)]
    out["warp_bgr"] = bgr; out["warp_gray"] = gray
    for i, (H, ds) in enumerate(zip(Hs, dsizes)):
        out[f"warp{i}_H"] = H; out[f"warp{i}_dsize"] = np.array(ds, np.int32)
        out[f"warp{i}_bgr"] = cv2.warpPerspective(bgr, H, ds)
        out[f"warp{i}_gray"] = cv2.warpPerspective(gray, H, ds)
        out[f"warp{i}_gray_inv"] = cv2.warpPerspective(gray, H, ds, flags=cv2.INTER_LINEAR | cv2.WARP_INVERSE_MAP)
    # ---- INTER_AREA prescale of u8/256 frames: fractional scales, integer scales 2 and 3, cut cells ----------------
    g2 = cv2.cvtColor(cv2.imread("/root/reference/doc/image.jpg"), cv2.COLOR_BGR2GRAY)
    frames = {"a": g2[40:300, 100:487], "b": g2[0:217, 0:385], "c": g2[100:318, 200:584]}   # 260x387, 217x385, 218x384
    cases = [("a", 300), ("a", 129), ("a", 386), ("b", 250), ("c", 192), ("c", 128), ("a", 500), ("b", 77)]
    out["area_cases"] = np.array([f"{k}:{m}" for k, m in cases])
    for k, v in frames.items():
        out[f"area_src_{k}"] = v.copy()
    for k, m in cases:
        f32 = frames[k].astype(np.float32) / np.float32(256.0)
        r, sc = prescale_cv(f32, m)
        out[f"area_{k}_{m}"] = r; out[f"area_{k}_{m}_scale"] = np.float32(sc)
    # a full-float frame (not u8-derived) at a fractional scale: products and sums round, the order matters
    ff = rng.random((131, 173)).astype(np.float32)
    out["area_float_src"] = ff
    out["area_float_100"], _ = prescale_cv(ff, 100)
    # ---- the whole demo flow on a BGR crop: cv2 image ops + the compiled reference ----------------------------------
    ref = Ref("canon")
    crop = cv2.imread("/root/reference/doc/image.jpg")[100:420, 200:680].copy()          # 320 x 480 x 3 (the ingest_u8 crop)
    for tag, max_size, refine in (("flow", 300, False), ("flow_full", 1200, False)):
        g = cv2.cvtColor(crop, cv2.COLOR_BGR2GRAY)
        f32 = g.astype(np.float32) * np.float32(1 / 256.0)
        small, sc = prescale_cv(f32, max_size)
        small = np.ascontiguousarray(small)
        ml = np.float32(max(small.shape)) / np.float32(100.0)
        lines = ref.find_line_segment_groups(small, float(ml), refine=refine, num_threads=-1, seed=0)
        for k in ("x1", "y1", "x2", "y2"):
            lines[k] = lines[k] / np.float32(sc)
        t = transform_to_dict(ref.compute_rectification_transform(lines, crop.shape[1], crop.shape[0], RectificationConfig(40.0, 1.5, 2, 2.0, 2)))
        H, ds = homography_from_corners_cv(t, 3.0)
        out["flow_src"] = crop; out[f"{tag}_max_size"] = np.float32(max_size); out[f"{tag}_lines"] = lines
        out[f"{tag}_corners"] = np.array([t["top_left"], t["top_right"], t["bottom_left"], t["bottom_right"]], np.float32)
        out[f"{tag}_H"] = H; out[f"{tag}_dsize"] = np.array(ds, np.int32)
        out[f"{tag}_warped"] = cv2.warpPerspective(crop, H, ds)
        print(tag, "scale", sc, "lines", len(lines), "dsize", ds)
    np.savez_compressed(os.path.join(GOLD, "warp_area.npz"), **out)
    print("warp_area.npz", os.path.getsize(os.path.join(GOLD, "warp_area.npz")))


if __name__ == "__main__":
    main()
