"""Time K1 (+K2) alone: one engine, n slots, CUDA-event stage timers.  Usage: python scripts/k1_time.py [slots] [reps]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api
from librectify_b200.synth import manhattan_frame
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
frames = torch.stack([manhattan_frame(i, 1920, 1080, device="cuda") for i in range(n)]).contiguous()
eng = api.Engine(0, n)
for _ in range(3):
    eng.rectify(frames, 19.2)
eng.profile(True)
for _ in range(reps):
    eng.rectify(frames, 19.2)
p = eng.profile_read()
px = n * 1920 * 1080
for k, b in (("grad", 17.0), ("nms", 6.0)):
    us = 1e3 * p["ms"][k] / reps
    print(f"{k}: {us:.1f} us per {n} frames, {b * px / us / 1e3:.0f} GB/s algorithmic, frac of 6526.8 = {b * px / us / 1e3 / 6526.8:.3f}")
