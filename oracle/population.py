"""TEST INFRASTRUCTURE ONLY — population parity: a set of frames through the CPU oracle, compared with results the CUDA path
produced for the same frames.  Used by tests/test_population_gpu.py and by bench.py's `parity_sample` (outside the timed
region).  The acceptance figure is the north star's: all four output corners within 0.5 px of the reference's."""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from .bind import Oracle, RectificationConfig, transform_to_dict

CORNERS = ("top_left", "top_right", "bottom_left", "bottom_right")


def oracle_results(frames: np.ndarray, min_length: float, cfg: RectificationConfig | None = None, threads: int | None = None):
    """[(segments or None, transform dict)] per frame; ctypes releases the GIL, so the frames run on all host cores"""
    orc = Oracle()
    cfg = cfg or RectificationConfig.default()
    h, w = frames.shape[1:]

    def one(i):
        segs = orc.find_line_segment_groups(np.ascontiguousarray(frames[i]), min_length, seed=0)
        n = 0 if segs is None else len(segs)
        s = segs if segs is not None else np.zeros(0, dtype=[("x1", "<f4"), ("y1", "<f4"), ("x2", "<f4"), ("y2", "<f4"), ("weight", "<f4"), ("err", "<f4"), ("group_id", "<i4")])
        return segs, transform_to_dict(orc.compute_rectification_transform(s, w, h, cfg)) if n else None

    with ThreadPoolExecutor(max_workers=threads or os.cpu_count() or 1) as ex:
        return list(ex.map(one, range(len(frames))))


def compare(gpu_lines, gpu_tf, ref, tol_px: float = 0.5) -> dict:
    """gpu_lines: list of SEG arrays (or None), gpu_tf: TF_DTYPE array, ref: oracle_results(...).  Returns the report."""
    n = len(ref)
    identical = within = both_empty = 0
    worst = 0.0
    for i in range(n):
        segs, tf = ref[i]
        gl = gpu_lines[i]
        g_empty = gl is None or len(gl) == 0
        if segs is None:
            ok = g_empty
            identical += ok; within += ok; both_empty += ok
            continue
        if g_empty:
            continue
        identical += (len(gl) == len(segs) and gl.tobytes() == segs.tobytes())
        d = 0.0
        for c in CORNERS:
            a = np.asarray(gpu_tf[i][c][:2], np.float64); b = np.asarray(tf[c][:2], np.float64)
            dd = float(np.max(np.abs(a - b))) if np.all(np.isfinite(a)) and np.all(np.isfinite(b)) else (0.0 if np.array_equal(a, b, equal_nan=True) else np.inf)
            d = max(d, dd)
        worst = max(worst, d)
        within += d <= tol_px
    return {"frames": n, "corners_within_tol_pct": 100.0 * within / max(n, 1), "tol_px": tol_px,
            "segments_and_groups_byte_identical_pct": 100.0 * identical / max(n, 1), "worst_corner_diff_px": worst,
            "frames_without_lines": both_empty}
