"""Frame sharding and the one-device runner used by the benchmark, the tests and batch users.

Frames are independent, so the path shards with no data-path collective: frame i of B goes to rank (or
device) floor(i * G / B) (contiguous blocks), and inside one device to one of a few engines (stream + workspace each).  The
sharding, the per-engine host threads and the gather live in the library (lrb_pool_*, csrc/pool_api.inc); this module is the
thin Python face of a one-device pool plus ``shard_bounds`` for multi-process launches (one rank per GPU under torchrun).
"""
from __future__ import annotations

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
        self.pool = api.Pool([device], n_engines, slots)
        self.engines = self.pool.engines
        self.device = device

    def close(self):
        self.pool.close()

    def profile(self, on=True):
        for e in self.engines:
            e.profile(on)

    def profile_read(self):
        tot = dict(ms={k: 0.0 for k in api.TIMERS}, groups=0, kernels=0, h2d_bytes=0, d2h_bytes=0)
        for e in self.engines:
            p = e.profile_read()
            for k in api.TIMERS:
                tot["ms"][k] += p["ms"][k]
            tot["groups"] += p["groups"]; tot["kernels"] += p["kernels"]
            tot["h2d_bytes"] += p["h2d_bytes"]; tot["d2h_bytes"] += p["d2h_bytes"]
        return tot

    def rectify(self, frames, min_length: float, cfg=None, u8: bool = False, steps: int = 1):
        """frames: [n, h, w] numpy (host; float32 or uint8) or torch CUDA tensor.  Returns (transforms[n], info[n]).

        steps > 1 streams the same batch through ``steps`` times (a benchmark's K steps, or K batches of a video): every
        engine goes through its shard of each step — input copy, kernels, read-back, corner solve — and moves on to the next
        step without waiting for the other engines.  The results of the last step are returned."""
        if isinstance(frames, np.ndarray):
            return self.pool.rectify_host(frames, min_length, cfg, passes=steps)
        return self.pool.rectify_device(frames, min_length, cfg, passes=steps)
