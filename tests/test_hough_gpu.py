"""Dormant Hough-voted estimator on the device (reference src/line_pencil.cpp:47-86,
src/estimator.h:82-96) through the C ABI (lrb_direct_estimate): weights and vanishing point bit-exact against the
compiled reference's outputs in tests/golden/hough.npz, and against the oracle on fresh index lists."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from librectify_b200 import api as a
    a.load_library()
    return a


def _segs(golden, name):
    g = golden(name)
    return g["segs"] if "segs" in g else g["lines"]


def test_direct_estimate_matches_reference_golden(api, golden):
    h = golden("hough")
    eng = api.Engine(0, 1)
    for case in h["cases"]:
        name, _ = str(case).split(":")
        segs = _segs(golden, name); idx = h[f"{case}:idx"]
        w, cell, vp = eng.direct_estimate(segs, idx)
        assert w.tobytes() == h[f"{case}:weights"].tobytes(), case
        assert vp.tobytes() == h[f"{case}:vp"].tobytes(), case
        assert 0 <= cell[0] < 65 and 0 <= cell[1] < 65
    eng.close()


def test_direct_estimate_matches_oracle_on_fresh_lists(api, golden):
    from oracle.bind import Oracle
    orc = Oracle()
    eng = api.Engine(0, 1)
    rng = np.random.default_rng(11)
    segs = _segs(golden, "doc_image")
    w, _, vp = eng.direct_estimate(segs)                             # idx = NULL: all lines
    allidx = np.arange(len(segs), dtype=np.int32)
    assert w.tobytes() == orc.get_weights(segs, allidx).tobytes() and vp.tobytes() == orc.direct_solve(segs, allidx).tobytes()
    for m in (2, 3, 50, 500):
        idx = rng.permutation(len(segs))[:m].astype(np.int32)        # unsorted lists too
        w, _, vp = eng.direct_estimate(segs, idx)
        assert w.tobytes() == orc.get_weights(segs, idx).tobytes(), m
        assert vp.tobytes() == orc.direct_solve(segs, idx).tobytes(), m
    bad = np.array([0, len(segs)], np.int32)
    with pytest.raises(api.LibrectifyError):
        eng.direct_estimate(segs, bad)
    eng.close()
