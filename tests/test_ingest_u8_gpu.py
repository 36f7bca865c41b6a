"""8-bit ingest on the device (reference src/autorectify.cpp:119 convertTo(CV_32F, 1/256.0) and :340
cvtColor BGR2GRAY), through the C ABI.  The fixture holds a BGR crop of the reference's bundled photo, cv2's own gray
conversion of it and the compiled reference's result on gray/256 (oracle/gen_golden.py): everything is compared bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from librectify_b200 import api as a
    a.load_library()
    return a


def test_gray_and_bgr_frames_match_the_float_path_and_the_reference(api, golden):
    g = golden("ingest_u8")
    ml = float(g["min_length"])
    eng = api.Engine(0, 1)
    _, lf = eng.detect(g["gray"].astype(np.float32) / np.float32(256.0), ml)
    assert lf[0].tobytes() == g["lines"].tobytes()
    _, lg = eng.detect_u8(g["gray"], ml)
    assert lg[0].tobytes() == g["lines"].tobytes()
    img = eng.fetch("dx")                                            # sanity: the workspace really came from the u8 path
    assert img.shape == g["gray"].shape
    _, lb = eng.detect_u8(g["bgr"], ml)                              # BGR2GRAY on the device == cv2's
    assert lb[0].tobytes() == g["lines"].tobytes()
    eng.close()


def test_u8_batches_host_and_device(api, golden):
    import torch
    g = golden("doc_image")
    gray = g["gray_u8"]; ml = float(g["min_length"])
    frames = np.stack([gray, gray[::-1].copy(), np.roll(gray, 17, axis=1), gray, gray[:, ::-1].copy()])
    e1 = api.Engine(0, 1); e2 = api.Engine(0, 2)
    ref = [e1.detect(f.astype(np.float32) / np.float32(256.0), ml)[1][0] for f in frames]
    _, host = e2.detect_u8(frames, ml)                               # 3 launch groups, the next one prefetched
    _, dev = e2.detect_u8(torch.from_numpy(frames).cuda(), ml)
    for i in range(5):
        assert host[i].tobytes() == ref[i].tobytes(), i
        assert dev[i].tobytes() == ref[i].tobytes(), i
    tf, _ = e2.rectify_u8(frames, ml)
    tf_f, _ = e2.rectify(np.stack([f.astype(np.float32) / np.float32(256.0) for f in frames]), ml)
    assert tf.tobytes() == tf_f.tobytes()
    e1.close(); e2.close()


def test_u8_argument_checks(api):
    eng = api.Engine(0, 1)
    bad = np.zeros((1, 32, 32, 2), np.uint8)                         # two channels: rejected, with a reason
    rc = eng.lib.lrb_detect_host_u8(eng.h, bad.ctypes.data, 1, 32, 32, 2, 64, 2048, 1.0, None, None, 0)
    assert rc != 0 and b"channels" in eng.lib.lrb_last_error()
    eng.close()
