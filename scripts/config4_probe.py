"""BASELINE config 4 (3840x2160 dense clutter, min_length 8): throughput (4 engines x 2 slots, 8 frames) and single-frame latency
with the per-stage breakdown."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api
from librectify_b200.runner import Runner
from librectify_b200.synth import manhattan_frame
if os.environ.get("LRB_LIB_PATH"):
    api.LIB_PATH = os.environ["LRB_LIB_PATH"]
W, H, L, ml, nf = 3840, 2160, 400, 8.0, 8
frames = torch.stack([manhattan_frame(100 + i, W, H, n_lines=L, device="cuda") for i in range(nf)]).contiguous()
cfg = api.make_config()
r = Runner(0, 4, 2)
for _ in range(2):
    r.rectify(frames, ml, cfg)
r.profile(True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(3):
    tf, info = r.rectify(frames, ml, cfg)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 3
p = r.profile_read(); r.close()
print("throughput %.1f images/s" % (nf / dt), {k: round(v / (3 * nf), 3) for k, v in p["ms"].items()}, "n_obs", info["n_obs"][0].tolist(), flush=True)
e1 = api.Engine(0, 1)
for _ in range(2):
    e1.rectify(frames[0], ml, cfg)
lat = []
for _ in range(6):
    torch.cuda.synchronize(); t0 = time.perf_counter(); e1.rectify(frames[0], ml, cfg); lat.append(1e3 * (time.perf_counter() - t0))
e1.profile(True)
for _ in range(4):
    e1.rectify(frames[0], ml, cfg)
p = e1.profile_read(); e1.close()
print("latency p50 %.2f ms" % np.median(lat), {k: round(v / 4, 3) for k, v in p["ms"].items()}, flush=True)
