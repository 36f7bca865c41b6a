"""lrb_fit_vanishing_point_noalloc: the allocation-free variant of fit_vanishing_point on the device (csrc/k_aux.cu; scratch in a
static __device__ array, no counterpart in the reference).  Called through the C ABI on caller-provided device memory and compared
with the existing entry point lrb_fit_vanishing_point_device — bit for bit, as every parity test of this suite — on calls in a row
that share the scratch array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _grouped_lines(rng, n, groups=(0, 1, 2, 3, 7, -1)):
    from librectify_b200 import api
    out = np.zeros(n, api.SEG_DTYPE)
    vps = np.array([rng.uniform(-3000, 5000, 2) for _ in groups])
    g = rng.integers(len(groups), size=n)
    a = rng.uniform(0, 1920, (n, 2)) * np.array([1, 1080 / 1920])
    d = vps[g] - a
    d = d / np.linalg.norm(d, axis=1, keepdims=True) + rng.normal(0, 0.01, (n, 2))
    b = a + d * rng.uniform(20, 300, (n, 1))
    out["x1"], out["y1"], out["x2"], out["y2"] = a[:, 0], a[:, 1], b[:, 0], b[:, 1]
    out["weight"] = rng.uniform(0.1, 1, n); out["err"] = rng.uniform(0, 1, n)
    out["group_id"] = np.array(groups)[g]
    return out


def _noalloc(lib, torch, lines, group, stream=None):
    """Device buffers come from torch (the caller's business); the entry point itself allocates nothing."""
    d_lines = torch.from_numpy(lines.view(np.uint8).copy()).cuda()
    d_out = torch.full((3,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    st = stream.cuda_stream if stream is not None else None
    rc = lib.lrb_fit_vanishing_point_noalloc(0, d_lines.data_ptr(), len(lines), int(group), d_out.data_ptr(), st)
    assert rc == 0, (rc, lib.lrb_last_error())
    (stream.synchronize() if stream is not None else torch.cuda.synchronize())
    return d_out.cpu().numpy()


def test_noalloc_equals_existing_entry_point_on_calls_in_a_row():
    import torch
    from librectify_b200 import api
    lib = api.load_library()
    eng = api.Engine(0, 1)
    try:
        rng = np.random.default_rng(2024)
        big, small = _grouped_lines(rng, 20000), _grouped_lines(rng, 37)
        full = _grouped_lines(rng, 65536)                        # exactly LRB_NOALLOC_MAX_LINES
        side = torch.cuda.Stream()
        # every call shares the one scratch array: long list, then a short one (stale model entries behind it), other groups,
        # the "all lines" cases (group <= 0, group without members), a non-default stream, the maximum size
        for lines, group, stream in ((big, 1, None), (small, 3, None), (small, 3, None), (big, 7, side), (small, -1, None),
                                     (big, 5, None), (full, 2, side), (small, 1, side), (big, 0, None)):
            want = np.float32(eng.fit_vanishing_point(lines, group))
            got = _noalloc(lib, torch, lines, group, stream)
            assert got.tobytes() == want.tobytes(), (len(lines), group, got, want)
            assert np.float32(api.fit_vanishing_point(lines, group)).tobytes() == got.tobytes()      # and the host version
    finally:
        eng.close()


def test_noalloc_rejects_what_does_not_fit():
    import torch
    from librectify_b200 import api
    lib = api.load_library()
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(3, dtype=torch.float32, device="cuda")
    assert lib.lrb_fit_vanishing_point_noalloc(0, d.data_ptr(), 65537, 1, out.data_ptr(), None) == -5
    assert b"LRB_NOALLOC_MAX_LINES" in lib.lrb_last_error() and lib.lrb_last_status() == -5
    assert lib.lrb_fit_vanishing_point_noalloc(0, d.data_ptr(), 0, 1, out.data_ptr(), None) == -1
    assert lib.lrb_fit_vanishing_point_noalloc(0, None, 5, 1, out.data_ptr(), None) == -1
    assert lib.lrb_fit_vanishing_point_noalloc(0, d.data_ptr(), 5, 1, None, None) == -1
