"""One small frame through the engine with the flood statistics printed (debugging aid).  Usage: python scripts/flood_probe.py [w h seed]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from librectify_b200 import api
from librectify_b200.synth import manhattan_frame
w = int(sys.argv[1]) if len(sys.argv) > 1 else 640
h = int(sys.argv[2]) if len(sys.argv) > 2 else 360
seed = int(sys.argv[3]) if len(sys.argv) > 3 else 7
img = manhattan_frame(seed, w, h).numpy()
eng = api.Engine(0, 1)
t0 = time.time()
info, lines = eng.detect(img, max(w, h) / 100.0)
print("time %.3f s" % (time.time() - t0), {k: info[0][k].tolist() for k in info.dtype.names}, flush=True)
t0 = time.time()
info, lines = eng.detect(img, max(w, h) / 100.0)
print("time %.3f s" % (time.time() - t0), flush=True)
