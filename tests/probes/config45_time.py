"""Throughput of BASELINE configs 4 (3840x2160 dense clutter, min_length 8) and 5 (4000x3000) on one B200 (device-resident
frames, 4 engines x 4 slots), and the reference's OpenMP CPU path on the same frames (oracle/_ref/librectify_ship.so)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from librectify_b200 import api
from librectify_b200.runner import Runner
from librectify_b200.synth import manhattan_frame
from oracle.bind import Ref, RectificationConfig
cores = os.cpu_count()
for name, W, H, L, ml, nf in (("config4 4K clutter", 3840, 2160, 400, 8.0, 16), ("config5 12MP", 4000, 3000, 40, 40.0, 16)):
    frames = torch.stack([manhattan_frame(100 + i, W, H, n_lines=L, device="cuda") for i in range(nf)]).contiguous()
    r = Runner(0, 4, 4)
    cfg = api.make_config()
    for _ in range(2):
        tf, info = r.rectify(frames, ml, cfg)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        tf, info = r.rectify(frames, ml, cfg)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / reps
    r.close()
    e1 = api.Engine(0, 1); e1.rectify(frames[:1], ml, cfg)
    torch.cuda.synchronize(); t0 = time.perf_counter(); e1.rectify(frames[:1], ml, cfg); torch.cuda.synchronize(); lat = time.perf_counter() - t0
    e1.close()
    ref = Ref("ship"); host = frames[:2].cpu().numpy()
    t0 = time.perf_counter()
    for i in range(2):
        segs = ref.find_line_segment_groups(host[i], ml, num_threads=cores, seed=0)
        ref.compute_rectification_transform(segs, W, H, RectificationConfig.default())
    cpu = (time.perf_counter() - t0) / 2
    print(f"{name}: {nf / dt:.1f} images/s (4 engines x 4 slots, {nf} frames), single-frame latency {1e3 * lat:.1f} ms, n_lines mean {info['n_lines'].mean():.0f}, "
          f"pairs/frame {1e4 * info['n_obs'].sum(axis=1).mean():.3g}; reference {cores} threads: {cpu:.2f} s/frame = {1 / cpu:.2f} images/s", flush=True)
