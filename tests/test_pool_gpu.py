"""The in-library multi-GPU batch entry (lrb_pool_*, csrc/pool_api.inc): frame i of B -> device floor(i * G / B), one host thread
per engine inside the library, results gathered in frame order — byte-identical to one engine doing every frame.

On a box with two or more GPUs the pool spans devices 0 and 1 in ONE process (which also covers the per-device kernel
attributes: every launcher's shared-memory opt-in must be made on each device); with a single GPU both shards land on it."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H, ML = 640, 360, 6.4


def _frames(n):
    from librectify_b200.synth import manhattan_frame
    return np.stack([manhattan_frame(300 + i, W, H, n_lines=30).numpy() for i in range(n)])


RESULT_FIELDS = ["n_lines", "n_seeds", "n_components", "n_pixels", "n_solves", "n_obs", "max_magnitude", "flood_fallback"]


def _same_info(a, b):
    """results identical; the flood statistics (rounds, floods per pass) depend on the launch-group size by design (window schedule)"""
    return all(np.array_equal(a[f], b[f]) for f in RESULT_FIELDS)


def test_pool_matches_single_engine():
    import torch
    from librectify_b200 import api
    frames = _frames(11)
    cfgs = np.concatenate([api.make_config(h_strategy=hs, v_strategy=vs) for hs, vs in ((2, 2), (0, 0), (3, 2))])
    one = api.Engine(0, 4)
    tf_ref, info_ref = one.rectify(frames, ML, cfgs)
    one.close()
    devices = [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]
    pool = api.Pool(devices, engines_per_device=2, slots=2)              # 11 frames over 2 x 2 engines x 2 slots: ragged shards
    tf, info = pool.rectify_host(frames, ML, cfgs)
    assert tf.tobytes() == tf_ref.tobytes() and _same_info(info, info_ref)
    tf2, _ = pool.rectify_host(frames, ML, cfgs, passes=3)
    assert tf2.tobytes() == tf_ref.tobytes()
    u8 = (frames * 256.0).clip(0, 255).astype(np.uint8)
    tf8, _ = pool.rectify_host(u8, ML, cfgs)
    one = api.Engine(0, 4)
    tf8_ref, _ = one.rectify_u8(u8, ML, cfgs)
    one.close()
    assert tf8.tobytes() == tf8_ref.tobytes()
    # device-resident shards: shard d lives on pool device d
    lo = len(frames) * 1 // 2
    shards = [torch.from_numpy(frames[:lo]).cuda(devices[0]), torch.from_numpy(frames[lo:]).cuda(devices[1])]
    tfd, infod = pool.rectify_device(shards, ML, cfgs)
    assert tfd.tobytes() == tf_ref.tobytes() and _same_info(infod, info_ref)
    pool.close()


def test_pool_rejects_bad_device():
    from librectify_b200 import api
    with pytest.raises(api.LibrectifyError):
        api.Pool([99], 1, 1)
