"""Dormant Hough-voted estimator (reference src/line_pencil.cpp:47-86 get_weights, src/estimator.h:82-96
DirectEstimator): the CPU restatement against the compiled reference's outputs (tests/golden/hough.npz, oracle/gen_golden.py
--hough-only) and, where oracle/_ref exists, against the compiled reference live.  Bit-exact."""
import numpy as np
import pytest

from oracle.bind import Oracle, Ref, ref_available


def _segs(golden, name):
    g = golden(name)
    return g["segs"] if "segs" in g else g["lines"]


def test_oracle_matches_golden(golden):
    h = golden("hough"); orc = Oracle()
    for case in h["cases"]:
        name, _ = str(case).split(":")
        segs = _segs(golden, name); idx = h[f"{case}:idx"]
        assert orc.get_weights(segs, idx).tobytes() == h[f"{case}:weights"].tobytes(), case
        assert orc.direct_solve(segs, idx).tobytes() == h[f"{case}:vp"].tobytes(), case


@pytest.mark.skipif(not ref_available("canon"), reason="compiled reference not built here")
def test_oracle_matches_reference_live(golden):
    ref = Ref("canon"); orc = Oracle()
    rng = np.random.default_rng(5)
    segs = _segs(golden, "synth_640x360")
    for m in (2, 17, 300, len(segs)):
        idx = np.sort(rng.choice(len(segs), m, replace=False)).astype(np.int32)
        assert orc.get_weights(segs, idx).tobytes() == ref.get_weights(segs, idx).tobytes(), m
        assert orc.direct_solve(segs, idx).tobytes() == ref.direct_solve(segs, idx).tobytes(), m
