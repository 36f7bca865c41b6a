"""Population parity (north star: >= 95 % of the frames rectified within 0.5 px of the reference's corners).

The benchmark's own frames (synthetic 1920x1080 Manhattan scenes, seeds 0..N-1, min_length 19.2, H+V RECTIFY) go through the
CUDA path (lrb_detect_device + lrb_rectify_device) and through the CPU oracle; the figure north_star asks for is asserted,
the stricter one — segments, order and group ids byte-identical — is asserted too and written to gpurun_out/ as a report."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N = int(os.environ.get("LRB_POPULATION", "256"))
W, H, MIN_LENGTH = 1920, 1080, 19.2


def test_population_1080p_against_oracle():
    import torch
    from librectify_b200 import api
    from librectify_b200.runner import Runner
    from librectify_b200.synth import manhattan_frame
    from oracle.population import compare, oracle_results

    frames = torch.stack([manhattan_frame(s, W, H, device="cuda") for s in range(N)]).contiguous()
    runner = Runner(0, 4, 16)
    tf, info = runner.rectify(frames, MIN_LENGTH, api.make_config())
    lines = []
    eng = runner.engines[0]
    for i0 in range(0, N, eng.slots):
        _, ll = eng.detect(frames[i0:i0 + eng.slots], MIN_LENGTH)
        lines += ll
    # the transforms come from the segments just compared: recompute them on the host from `lines` and require identity
    for i in (0, N // 2, N - 1):
        assert api.tform_vec(api.compute_rectification_transform(lines[i], W, H)).tobytes() == api.tform_vec(tf[i]).tobytes()
    assert [len(l) for l in lines] == info["n_lines"].tolist()
    ref = oracle_results(frames.cpu().numpy(), MIN_LENGTH)
    rep = compare(lines, tf, ref)
    rep["seeds"] = f"0..{N - 1}"
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/population_1080p.json", "w") as f:
        json.dump(rep, f)
    print(rep)
    runner.close()
    assert rep["corners_within_tol_pct"] >= 95.0, rep                        # the north star's acceptance figure
    assert rep["segments_and_groups_byte_identical_pct"] >= 95.0, rep         # what this repo actually aims at: identity
