"""Frame sharding and the multi-engine runner used by the benchmark and by batch users.

Frames are independent, so the path shards with no data-path collective: frame i of B goes to rank
floor(i * G / B) (contiguous blocks), and inside one rank to one of a few engines (stream + workspace each) driven by
host threads — ctypes releases the GIL for the duration of a call, so the engines overlap on the GPU and hide each
other's latency-bound stages.  Results are gathered on the host; NCCL is not on the image path.
"""
from __future__ import annotations

import threading

import numpy as np

from . import api


def shard_bounds(n_items: int, n_parts: int, part: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of ``part`` when n_items are split over n_parts (sizes differ by at most one)."""
    assert 0 <= part < n_parts
    lo = (n_items * part) // n_parts
    hi = (n_items * (part + 1)) // n_parts
    return lo, hi


class Runner:
    """``n_engines`` engines with ``slots`` frame slots each on one device; rectify() splits a batch across them."""

    def __init__(self, device: int = 0, n_engines: int = 4, slots: int = 8):
        self.engines = [api.Engine(device, slots) for _ in range(n_engines)]
        self.device = device

    def close(self):
        for e in self.engines:
            e.close()

    def profile(self, on=True):
        for e in self.engines:
            e.profile(on)

    def profile_read(self):
        tot = dict(ms={k: 0.0 for k in api.TIMERS}, groups=0, kernels=0)
        for e in self.engines:
            p = e.profile_read()
            for k in api.TIMERS:
                tot["ms"][k] += p["ms"][k]
            tot["groups"] += p["groups"]; tot["kernels"] += p["kernels"]
        return tot

    def rectify(self, frames, min_length: float, cfg=None, u8: bool = False, steps: int = 1):
        """frames: [n, h, w] numpy (host) or torch CUDA tensor.  Returns (transforms[n], info[n]).

        steps > 1 streams the same batch through ``steps`` times (a benchmark's K steps, or K batches of a video): every
        engine thread goes through its shard of each step — input copy, kernels, read-back, corner solve — and moves on to
        the next step without waiting for the other engines, so one engine's host->device copy overlaps the others'
        kernels instead of all engines copying at the start of every step.  The results of the last step are returned."""
        n = frames.shape[0]
        k = len(self.engines)
        tf = np.zeros(n, api.TF_DTYPE); info = np.zeros(n, api.INFO_DTYPE)
        errs = []

        def work(j):
            lo, hi = shard_bounds(n, k, j)
            if hi <= lo:
                return
            try:
                eng = self.engines[j]
                for _ in range(max(1, steps)):
                    t, i = (eng.rectify_u8 if u8 else eng.rectify)(frames[lo:hi], min_length, cfg)
                tf[lo:hi] = t; info[lo:hi] = i
            except Exception as ex:  # surfaced to the caller below
                errs.append(ex)

        if k == 1:
            work(0)
        else:
            th = [threading.Thread(target=work, args=(j,)) for j in range(k)]
            for t in th:
                t.start()
            for t in th:
                t.join()
        if errs:
            raise errs[0]
        return tf, info
