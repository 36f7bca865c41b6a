"""TEST INFRASTRUCTURE ONLY — writes tests/golden/*.npz from the COMPILED, UNMODIFIED reference.

Run in the build container (needs /root/reference and `make -C oracle ref`):

    python -m oracle.gen_golden

Every array below is produced by oracle/_ref/librectify_canon.so (the reference's own sources, strict-IEEE flags,
fixed-seed shim — see oracle/Makefile); nothing comes from the restatement or from the CUDA path.  The fixtures are
what pins the restatement (tests/test_oracle_golden.py) and the CUDA path (tests/test_gpu_*.py) on machines where
/root/reference does not exist (the GPU box).

Cases
  kat.npz            sampler KATs (reference choice_knuth + std::mt19937), reference src/test.cpp known answers,
                     struct sizes, 5x5 taps
  synth_320x240.npz  full-float synthetic frame: every detector stage + grouping + all 16 strategy combinations
  synth_640x360.npz  second synthetic frame (another seed), final outputs + stage digests
  doc_image.npz      reference doc/image.jpg (BASELINE config 1) as cv2 BGR2GRAY u8: final segments + transform
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.bind import Ref, RectificationConfig, SEG_DTYPE, transform_to_dict  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def digest(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def tform_vec(t) -> np.ndarray:
    d = transform_to_dict(t)
    return np.array([d["width"], d["height"], *d["top_left"], *d["top_right"], *d["bottom_left"], *d["bottom_right"],
                     *d["horizontal_vp"], *d["vertical_vp"]], dtype=np.float64)


def pairs_hash(pairs: np.ndarray) -> str:
    """FNV-1a-style hash over index words, as defined in SURVEY.md Appendix C."""
    h = 1469598103934665603
    for v in pairs.reshape(-1).tolist():
        h = ((h ^ int(v)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


def frame_case(ref: Ref, img: np.ndarray, min_length: float, full_stages: bool, cfg=None) -> dict:
    h, w = img.shape
    st = ref.detect_stages(img)
    out = dict(image=img, min_length=np.float32(min_length))
    for k in ("dx", "dy", "mag", "bin", "labels"):
        out["sha_" + k] = digest(st[k])
    out["sha_grad8"] = digest(st["grad8"])
    out["seeds"] = st["seeds"]
    out["comp_sizes"] = st["comp_sizes"]
    out["raw_segs"] = st["segs"]
    if full_stages:
        out["bin_u8"] = st["bin"].astype(np.uint8)
        out["labels"] = st["labels"]
        out["mag"] = st["mag"]
    segs = ref.find_line_segment_groups(img, min_length, seed=0)
    out["segs"] = segs if segs is not None else np.zeros(0, SEG_DTYPE)
    if segs is not None:
        m = ref.pencil_model(segs)
        for k in ("bb", "center", "scale", "h", "anchor", "dir", "len"):
            out["model_" + k] = m[k]
        tf = []
        for hs in range(4):
            for vs in range(4):
                c = cfg(hs, vs) if cfg else RectificationConfig(40.0, 1.5, vs, 1.5, hs)
                tf.append(tform_vec(ref.compute_rectification_transform(segs, w, h, c)))
        out["tforms"] = np.stack(tf)          # index = h_strategy*4 + v_strategy
    return out


def refine_case(ref: Ref) -> dict:
    """`refine` (reference src/line_detector.cpp:340-444): raw segments, their merge, and the final grouped segments"""
    from librectify_b200.synth import manhattan_frame
    img = manhattan_frame(11, 480, 270, n_lines=25).numpy()
    st = ref.detect_stages(img, want_grad8=False)
    return dict(image=img, min_length=np.float32(5.0), raw=st["segs"], merged=ref.refine(st["segs"]),
                lines_refined=ref.find_line_segment_groups(img, 5.0, refine=True, num_threads=-1, seed=0),
                lines_plain=ref.find_line_segment_groups(img, 5.0, refine=False, num_threads=-1, seed=0))


def ingest_case(ref: Ref) -> dict:
    """8-bit ingest: a BGR crop of the bundled photo, OpenCV's BGR2GRAY of it (cv2, this container), and the
    reference's result on gray/256 — what src/autorectify.cpp:119,340 feed the library"""
    import cv2
    bgr = cv2.imread("/root/reference/doc/image.jpg")[100:420, 200:680].copy()
    gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
    img = gray.astype(np.float32) / np.float32(256.0)
    return dict(bgr=bgr, gray=gray, min_length=np.float32(5.0),
                lines=ref.find_line_segment_groups(img, 5.0, refine=False, num_threads=-1, seed=0))


def hough_case(ref: Ref) -> dict:
    """dormant estimator: LinePencilModel::get_weights + DirectEstimator::solve of the compiled reference
    (templates instantiated by oracle/ref_harness.cpp) on the segments of the committed fixtures"""
    out = {}
    names = []
    for name in ("doc_image", "synth_320x240", "synth_640x360", "ingest_u8"):
        g = np.load(os.path.join(GOLD, name + ".npz"))
        segs = g["segs"] if "segs" in g.files else g["lines"]
        n = len(segs)
        lists = {"all": np.arange(n, dtype=np.int32), "third": np.arange(0, n, 3, dtype=np.int32),
                 "rest": np.nonzero(segs["group_id"] != 0)[0].astype(np.int32), "few": np.array([5, 1, 9, 2], np.int32)}
        for k, idx in lists.items():
            tag = f"{name}:{k}"; names.append(tag)
            out[tag + ":idx"] = idx
            out[tag + ":weights"] = ref.get_weights(segs, idx)
            out[tag + ":vp"] = ref.direct_solve(segs, idx)
    out["cases"] = np.array(names)
    return out


PROSAC_TOL = float(np.float32(1.0) - np.float32(np.cos(np.float32(2.0) / np.float32(180) * np.float32(np.pi))))   # 1 - cos(2 deg), line_pencil.cpp:143-146


def prosac_case(ref: Ref) -> dict:
    """dormant estimator: PROSAC_Estimator<LinePencilModel>::solve of the compiled reference (reference
    src/prosac.h:104-299, instantiated by oracle/ref_harness.cpp, generator seeded through the random_device shim) on the
    segments of the committed fixtures, several index lists and seeds"""
    out = {}
    names = []
    for name in ("doc_image", "synth_320x240", "synth_640x360", "ingest_u8"):
        g = np.load(os.path.join(GOLD, name + ".npz"))
        segs = g["segs"] if "segs" in g.files else g["lines"]
        n = len(segs)
        lists = {"all": np.arange(n, dtype=np.int32), "third": np.arange(0, n, 3, dtype=np.int32),
                 "rest": np.nonzero(segs["group_id"] != 0)[0].astype(np.int32), "dozen": np.arange(40, 52, dtype=np.int32)}
        for k, idx in lists.items():
            for seed in (0, 77):
                tag = f"{name}:{k}:{seed}"; names.append(tag)
                out[tag + ":idx"] = idx
                out[tag + ":vp"] = ref.prosac_solve(segs, idx, PROSAC_TOL, seed)
    out["cases"] = np.array(names)
    out["tol"] = np.float32(PROSAC_TOL)
    return out


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = Ref("canon")
    if "--prosac-only" in sys.argv:                       # added after the other fixtures were committed: leave those untouched
        np.savez_compressed(os.path.join(GOLD, "prosac.npz"), **prosac_case(ref))
        return
    if "--hough-only" in sys.argv:                        # added after the other fixtures were committed: leave those untouched
        np.savez_compressed(os.path.join(GOLD, "hough.npz"), **hough_case(ref))
        return
    from librectify_b200.synth import manhattan_frame
    if "--ingest-only" in sys.argv:
        np.savez_compressed(os.path.join(GOLD, "ingest_u8.npz"), **ingest_case(ref))
        return
    if "--refine-only" in sys.argv:                       # added after the other fixtures were committed: leave those untouched
        np.savez_compressed(os.path.join(GOLD, "refine.npz"), **refine_case(ref))
        return

    # ---- KATs ------------------------------------------------------------------------------------------------
    kat = {}
    kat["sizeof"] = np.array([ref.sizeof(i) for i in range(4)], np.int32)
    kat["taps_x"] = ref.gauss_deriv_kernel(True)
    kat["taps_y"] = ref.gauss_deriv_kernel(False)
    p10, d10, n10 = ref.choice_knuth(0, 10, 5)
    kat["knuth_n10"] = p10
    for n in (940, 1570):
        p, d, nx = ref.choice_knuth(0, n, 10000)
        kat[f"knuth_n{n}_head"] = p[:64].copy()
        kat[f"knuth_n{n}_tail"] = p[-1].copy()
        kat[f"knuth_n{n}_draws"] = np.int64(d)
        kat[f"knuth_n{n}_next"] = np.uint32(nx)
        kat[f"knuth_n{n}_hash"] = pairs_hash(p)
    p, d, nx = ref.choice_knuth(12345, 37, 2000)
    kat["knuth_s12345_n37"] = p
    kat["knuth_s12345_n37_next"] = np.uint32(nx)
    # reference src/test.cpp
    ls = np.array([(0, 0, 10, 0, 1, 0, 10), (10, 0, 8, 5, 1, 0, 1), (8, 5, 2, 5, 1, 0, 10), (2, 5, 0, 0, 1, 0, 1)], SEG_DTYPE)
    kat["test0_lines"] = ls
    kat["test0_vp1"] = np.array(ref.fit_vanishing_point(ls, 1), np.float32)
    kat["test0_vp2"] = np.array(ref.fit_vanishing_point(ls, 10), np.float32)
    kat["test0_vp3"] = np.array(ref.fit_vanishing_point(ls, -1), np.float32)
    new = np.array([(5, 1, 5, 4, 1, 0, -1)], SEG_DTYPE)
    kat["test0_assigned"] = ref.assign_to_group(ls, new, 10.0)
    kat["from_vp"] = tform_vec(ref.compute_rectification_transform_from_vp(1920, 1080, (3000.0, 500.0, 1.0), (900.0, -4000.0, 1.0)))
    kat["from_vp_ideal"] = tform_vec(ref.compute_rectification_transform_from_vp(640, 480, (1.0, 0.0, 0.0), (0.0, 1.0, 0.0)))
    np.savez_compressed(os.path.join(GOLD, "kat.npz"), **kat)

    # ---- frames ----------------------------------------------------------------------------------------------
    img = manhattan_frame(3, 320, 240, n_lines=20).numpy()
    np.savez_compressed(os.path.join(GOLD, "synth_320x240.npz"), **frame_case(ref, img, 5.0, True))
    img = manhattan_frame(7, 640, 360, n_lines=30).numpy()
    np.savez_compressed(os.path.join(GOLD, "synth_640x360.npz"), **frame_case(ref, img, 6.4, False))

    doc = "/root/reference/doc/image.jpg"
    if os.path.exists(doc):
        import cv2
        g = cv2.cvtColor(cv2.imread(doc), cv2.COLOR_BGR2GRAY)
        imgf = (g.astype(np.float32) / np.float32(256.0))
        case = frame_case(ref, imgf, 10.0, False, cfg=lambda hs, vs: RectificationConfig(40.0, 1.5, vs, 2.0, hs))
        case.pop("image")
        case["gray_u8"] = g
        np.savez_compressed(os.path.join(GOLD, "doc_image.npz"), **case)
    np.savez_compressed(os.path.join(GOLD, "refine.npz"), **refine_case(ref))
    np.savez_compressed(os.path.join(GOLD, "ingest_u8.npz"), **ingest_case(ref))
    np.savez_compressed(os.path.join(GOLD, "hough.npz"), **hough_case(ref))
    np.savez_compressed(os.path.join(GOLD, "prosac.npz"), **prosac_case(ref))
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
