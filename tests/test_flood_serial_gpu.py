"""The serial fallback of the flood (k_flood_serial: the reference's region growing verbatim, reference src/filter.cpp:110-153,
src/line_detector.cpp:92-122) is the safety net behind the speculative rounds; real frames never need it, so it is forced here
(LRB_FORCE_SERIAL: one speculative round, the rest of the seeds by the serial kernel) and compared with the oracle like every other
path — segments, order and group ids byte for byte — alone and inside a batch."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def forced_serial():
    os.environ["LRB_FORCE_SERIAL"] = "1"
    yield
    os.environ.pop("LRB_FORCE_SERIAL", None)


@pytest.mark.parametrize("size,seed", [((333, 201), 11), ((640, 360), 7), ((1280, 720), 5)])
def test_serial_fallback_equals_oracle(forced_serial, oracle, size, seed):
    from librectify_b200 import api
    from librectify_b200.synth import manhattan_frame
    w, h = size
    img = manhattan_frame(seed, w, h).numpy()
    ml = max(w, h) / 100.0
    eng = api.Engine(0, 1)
    try:
        info, lines = eng.detect(img, ml, lines_cap=20000)
        assert info["flood_fallback"][0] == 1                                # the serial kernel did run
        ref = oracle.find_line_segment_groups(img, ml, seed=0)
        assert ref is not None and len(ref) == len(lines[0])
        assert lines[0].tobytes() == ref.tobytes()
    finally:
        eng.close()
    os.environ.pop("LRB_FORCE_SERIAL", None)                                 # and the speculative path gives the same bytes
    eng = api.Engine(0, 1)
    try:
        info2, lines2 = eng.detect(img, ml, lines_cap=20000)
        assert info2["flood_fallback"][0] == 0
        assert lines2[0].tobytes() == lines[0].tobytes()
    finally:
        eng.close()


def test_serial_fallback_in_a_batch(forced_serial, oracle):
    from librectify_b200 import api
    from librectify_b200.synth import manhattan_frame
    frames = np.stack([manhattan_frame(20 + i, 517, 389).numpy() for i in range(5)])
    eng = api.Engine(0, 4)                                                   # 5 frames on 4 slots: two launch groups
    try:
        info, lines = eng.detect(frames, 5.17, lines_cap=20000)
        for i in range(5):
            ref = oracle.find_line_segment_groups(frames[i], 5.17, seed=0)
            assert info["flood_fallback"][i] == 1
            assert lines[i].tobytes() == ref.tobytes(), i
    finally:
        eng.close()
