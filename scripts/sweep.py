"""Throughput of the device-resident path for a few runner shapes.  Usage: python scripts/sweep.py E:S:F [E:S:F ...]
(E engines x S slots, F frames per step).  Prints images/s, host-e2e images/s and the per-stage milliseconds per frame."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api  # noqa: E402
from librectify_b200.runner import Runner  # noqa: E402
from librectify_b200.synth import manhattan_frame  # noqa: E402

if os.environ.get("LRB_LIB_PATH"):           # A/B runs inside one gpurun call: another build of the library
    api.LIB_PATH = os.environ["LRB_LIB_PATH"]
W, H, ML = 1920, 1080, 19.2
shapes = [tuple(int(x) for x in a.split(":")) for a in sys.argv[1:]] or [(4, 8, 64)]
fmax = max(f for _, _, f in shapes)
frames = torch.stack([manhattan_frame(i, W, H, device="cuda") for i in range(fmax)]).contiguous()
host = torch.empty((fmax, H, W), dtype=torch.float32, pin_memory=True); host.copy_(frames)
cfg = api.make_config()
for e, s, f in shapes:
    r = Runner(0, e, s)
    for _ in range(3):
        r.rectify(frames[:f], ML, cfg)
    r.profile(True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3):
        tf, info = r.rectify(frames[:f], ML, cfg)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    p = r.profile_read(); r.profile(False)
    r.rectify(host[:f].numpy(), ML, cfg)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3):
        r.rectify(host[:f].numpy(), ML, cfg)
    torch.cuda.synchronize(); dth = time.perf_counter() - t0
    st = {k: round(v / (3 * f), 4) for k, v in p["ms"].items()}
    print(e, s, f, round(3 * f / dt), round(3 * f / dth), st, "fallbacks", int(info["flood_fallback"].sum()), flush=True)
    r.close()
