"""Dormant PROSAC estimator on the device (reference src/prosac.h:104-299) through the C ABI
(lrb_prosac_estimate): vanishing point bit-exact against the compiled reference's outputs in tests/golden/prosac.npz, and the
whole trace (samples drawn, inlier counts, growth and termination state) against the oracle on fresh index lists and seeds."""
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


def test_prosac_matches_reference_golden(api, golden):
    g = golden("prosac"); tol = float(g["tol"])
    for seed in (0, 77):
        api.set_seed(seed)
        eng = api.Engine(0, 1)
        for case in g["cases"]:
            name, _, s = str(case).split(":")
            if int(s) != seed:
                continue
            vp, tr = eng.prosac_estimate(_segs(golden, name), g[f"{case}:idx"], tol)
            assert vp.tobytes() == g[f"{case}:vp"].tobytes(), case
        eng.close()
    api.set_seed(0)


def test_prosac_trace_matches_oracle(api, golden):
    from oracle.bind import Oracle
    orc = Oracle()
    g = golden("prosac"); tol = float(g["tol"])
    rng = np.random.default_rng(21)
    segs = _segs(golden, "doc_image")
    for seed in (0, 5):
        api.set_seed(seed)
        eng = api.Engine(0, 1)
        for m in (12, 13, 60, 400, len(segs)):
            idx = rng.permutation(len(segs))[:m].astype(np.int32)        # unsorted lists too
            vp, tr = eng.prosac_estimate(segs, idx, tol)
            want, wtr = orc.prosac_solve(segs, idx, tol, seed, trace=True)
            for k in ("iterations", "n_final", "n_star", "I_N_best"):
                assert tr[k] == wtr[k], (m, seed, k, tr, wtr)
            assert np.array_equal(tr["samples"], wtr["samples"]) and np.array_equal(tr["I_N"], wtr["I_N"]), (m, seed)
            assert vp.tobytes() == want.tobytes(), (m, seed)
        with pytest.raises(api.LibrectifyError):
            eng.prosac_estimate(segs, np.arange(5, dtype=np.int32), tol)  # the reference reads past a list this short
        eng.close()
    api.set_seed(0)
