"""Debugging aid: the golden PROSAC cases one by one through lrb_prosac_estimate (prints the case before running it)."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from librectify_b200 import api
G = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
g = dict(np.load(os.path.join(G, "prosac.npz")))
eng = api.Engine(0, 1)
for case in g["cases"]:
    name, k, seed = str(case).split(":")
    if int(seed) != 0:
        continue
    d = dict(np.load(os.path.join(G, name + ".npz")))
    segs = d["segs"] if "segs" in d else d["lines"]
    idx = g[case + ":idx"]
    print(case, len(segs), len(idx), flush=True)
    vp, tr = eng.prosac_estimate(segs, idx, float(g["tol"]))
    print("   ", vp.tobytes() == g[case + ":vp"].tobytes(), tr["iterations"], tr["n_final"], tr["I_N"].tolist(), flush=True)
