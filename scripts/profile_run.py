"""Short, fixed run for ncu: one engine, 8 slots, 8 synthetic 1080p frames, a few groups (see profiles/README.md)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api  # noqa: E402
from librectify_b200.synth import manhattan_frame  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
frames = torch.stack([manhattan_frame(i, 1920, 1080, device="cuda") for i in range(n)]).contiguous()
eng = api.Engine(0, n)
for _ in range(reps):
    tf, info = eng.rectify(frames, 19.2)
print("n_lines", info["n_lines"].tolist(), "rounds", info["flood_rounds"].tolist())
