"""fit_vanishing_point / assign_to_group on the device (csrc/k_aux.cu), bit-identical to the host versions
(transform_host.cpp, reference src/interface.cpp:211-265) and checked against the oracle and the reference's known answers."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from librectify_b200 import api
    e = api.Engine(0, 1)
    yield e
    e.close()


def _random_grouped_lines(rng, n, groups=(0, 1, 2, 3, 7, -1)):
    from librectify_b200 import api
    vps = {g: rng.uniform(-3000, 5000, 2) for g in groups}
    out = np.zeros(n, api.SEG_DTYPE)
    for i in range(n):
        g = groups[int(rng.integers(len(groups)))]
        a = rng.uniform(0, 1920, 2) * np.array([1, 1080 / 1920])
        d = vps[g] - a; d = d / np.linalg.norm(d)
        d = d + rng.normal(0, 0.01, 2)
        b = a + d * rng.uniform(20, 300)
        out[i] = (a[0], a[1], b[0], b[1], rng.uniform(0.1, 1), rng.uniform(0, 1), g)
    return out


def test_known_answers_of_the_reference_smoke_program(eng, golden):
    from librectify_b200 import api
    kat = golden("kat")
    ls = kat["test0_lines"]
    for g, key in ((1, "test0_vp1"), (10, "test0_vp2"), (-1, "test0_vp3")):
        assert np.array_equal(np.float32(eng.fit_vanishing_point(ls, g)), kat[key])
    new = np.array([(5, 1, 5, 4, 1, 0, -1)], dtype=ls.dtype)
    assert eng.assign_to_group(ls, new, 10.0)["group_id"][0] == 1


@pytest.mark.parametrize("n,n_new", [(7, 3), (500, 1000), (20000, 200000)])
def test_device_equals_host_and_oracle(eng, oracle, n, n_new):
    from librectify_b200 import api
    rng = np.random.default_rng(n)
    lines = _random_grouped_lines(rng, n)
    fresh = _random_grouped_lines(rng, n_new)
    fresh["group_id"] = -1
    for g in (0, 1, 3, 7, 5, -1):                                # 5: no member -> every line; 0 and -1: every line (reference `g > 0`)
        assert np.float32(eng.fit_vanishing_point(lines, g)).tobytes() == np.float32(api.fit_vanishing_point(lines, g)).tobytes(), g
    for tol in (2.0, 10.0, 45.0):
        dev = eng.assign_to_group(lines, fresh, tol)
        host = api.assign_to_group(lines, fresh, tol)
        assert dev.tobytes() == host.tobytes()
        if n_new <= 1000:
            assert np.array_equal(oracle.assign_to_group(lines, fresh, tol)["group_id"], dev["group_id"])
    for g in (1, 7, -1):
        assert np.allclose(eng.fit_vanishing_point(lines, g), oracle.fit_vanishing_point(lines, g), rtol=1e-4, atol=1e-2)
    assert (eng.assign_to_group(lines, fresh, 10.0)["group_id"] >= 0).sum() > 0
