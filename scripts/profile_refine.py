import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from librectify_b200 import api
from librectify_b200.synth import manhattan_frame
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
frames = torch.stack([manhattan_frame(i, 1920, 1080, device="cuda") for i in range(n)]).contiguous()
eng = api.Engine(0, n); eng.set_refine(True)
for _ in range(2):
    tf, info = eng.rectify(frames, 19.2)
print(info["n_lines"].tolist())
