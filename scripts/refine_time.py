"""Cost of the refine post-process: one engine, n slots, stage timers with refine off / on."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api
from librectify_b200.synth import manhattan_frame
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
frames = torch.stack([manhattan_frame(i, 1920, 1080, device="cuda") for i in range(n)]).contiguous()
eng = api.Engine(0, n)
for on in (False, True):
    eng.set_refine(on)
    for _ in range(3):
        eng.rectify(frames, 19.2)
    eng.profile(True)
    for _ in range(5):
        tf, info = eng.rectify(frames, 19.2)
    p = eng.profile_read(); eng.profile(False)
    print("refine", on, "model stage %.1f us/frame" % (1e3 * p["ms"]["model"] / (5 * n)), "lines", info["n_lines"][:6].tolist(), "comps", info["n_components"][:3].tolist())
