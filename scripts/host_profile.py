"""Per-stage device time of the host-input path against the device-resident path (4 engines x 32 slots, 256 frames)."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from librectify_b200 import api
from librectify_b200.runner import Runner
from librectify_b200.synth import manhattan_frame
W, H, ML, f = 1920, 1080, 19.2, 256
frames = torch.stack([manhattan_frame(i, W, H, device="cuda") for i in range(f)]).contiguous()
host = torch.empty((f, H, W), dtype=torch.float32, pin_memory=True); host.copy_(frames)
cfg = api.make_config()
r = Runner(0, 4, 32)
for name, src in (("device", frames), ("host", host.numpy()), ("device", frames), ("host", host.numpy())):
    for _ in range(2):
        r.rectify(src, ML, cfg)
    r.profile(True)
    for e in r.engines: e.lib.lrb_profile_reset(e.h)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(4):
        r.rectify(src, ML, cfg)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    p = r.profile_read(); r.profile(False)
    st = {k: round(v / (4 * f), 4) for k, v in p["ms"].items()}
    print(name, round(4 * f / dt), st, "sum", round(sum(st.values()), 3), flush=True)
r.close()
