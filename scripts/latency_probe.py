"""Single-frame latency (BASELINE config 2) with a per-stage breakdown, for a few window schedules of the flood.
Usage: python scripts/latency_probe.py [DIV ...]   (LRB_WINDOW_DIV values; default: the library's own schedule only)"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api  # noqa: E402
from librectify_b200.synth import manhattan_frame  # noqa: E402

cfg = api.make_config()
frames = [manhattan_frame(s, 1920, 1080, device="cuda").contiguous() for s in (1, 2, 3)]
for div in (sys.argv[1:] or ["default"]):
    if div == "default":
        os.environ.pop("LRB_WINDOW_DIV", None)
    else:
        os.environ["LRB_WINDOW_DIV"] = div
    one = api.Engine(0, 1)
    for f in frames:
        for _ in range(5):
            one.rectify(f, 19.2, cfg)
    lat = []
    for i in range(45):
        f = frames[i % 3]
        torch.cuda.synchronize(); t0 = time.perf_counter()
        tf, info = one.rectify(f, 19.2, cfg)
        lat.append(1e3 * (time.perf_counter() - t0))
    one.profile(True)
    for i in range(12):
        tf, info = one.rectify(frames[i % 3], 19.2, cfg)
    p = one.profile_read(); one.profile(False)
    print("div", div, "p50 %.3f p90 %.3f ms" % (np.median(lat), np.percentile(lat, 90)), "rounds", int(info["flood_rounds"][0]),
          {k: round(v / 12, 3) for k, v in p["ms"].items()}, "kernels", p["kernels"] / 12, flush=True)
    one.close()
