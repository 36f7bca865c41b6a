"""Short, fixed run for ncu: python scripts/profile_run.py [frames=8] [reps=3] [W H n_lines min_length] — one engine, `frames` slots
of synthetic frames (default 1080p; `1 3 3840 2160 400 8.0` is one BASELINE config 4 frame)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api  # noqa: E402
from librectify_b200.synth import manhattan_frame  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
W, H, L, ml = (int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), float(sys.argv[6])) if len(sys.argv) > 6 else (1920, 1080, 40, 19.2)
frames = torch.stack([manhattan_frame((100 if L != 40 else 0) + i, W, H, n_lines=L, device="cuda") for i in range(n)]).contiguous()
eng = api.Engine(0, n)
for _ in range(reps):
    tf, info = eng.rectify(frames, ml)
print("n_lines", info["n_lines"].tolist(), "rounds", info["flood_rounds"].tolist())
