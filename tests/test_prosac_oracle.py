"""Dormant PROSAC estimator (reference src/prosac.h:104-299, never instantiated by the reference's own build):
the CPU restatement against the compiled reference's outputs (tests/golden/prosac.npz, oracle/gen_golden.py --prosac-only) and,
where oracle/_ref exists, against the compiled reference live.  Bit-exact."""
import numpy as np
import pytest

from oracle.bind import Oracle, Ref, ref_available


def _segs(golden, name):
    g = golden(name)
    return g["segs"] if "segs" in g else g["lines"]


def test_oracle_matches_golden(golden):
    g = golden("prosac"); orc = Oracle()
    tol = float(g["tol"])
    for case in g["cases"]:
        name, _, seed = str(case).split(":")
        vp, tr = orc.prosac_solve(_segs(golden, name), g[f"{case}:idx"], tol, int(seed), trace=True)
        assert vp.tobytes() == g[f"{case}:vp"].tobytes(), case
        assert 1 <= tr["iterations"] <= 9                       # T_N = niter_RANSAC(0.9, 0.5, 2) = 9 (prosac.h:117)


@pytest.mark.skipif(not ref_available("canon"), reason="compiled reference not built here")
def test_oracle_matches_reference_live(golden):
    ref = Ref("canon"); orc = Oracle()
    g = golden("prosac"); tol = float(g["tol"])
    rng = np.random.default_rng(9)
    segs = _segs(golden, "synth_640x360")
    for m in (12, 40, 300, len(segs)):
        idx = np.sort(rng.choice(len(segs), m, replace=False)).astype(np.int32)
        for seed in (0, 3, 1234):
            assert orc.prosac_solve(segs, idx, tol, seed).tobytes() == ref.prosac_solve(segs, idx, tol, seed).tobytes(), (m, seed)
