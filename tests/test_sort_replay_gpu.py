"""Tie order of the seeds: the CUDA replay of libstdc++'s introsort (librectify_b200/csrc/k3_replay.cu) against std::sort itself.

The reference sorts its peaks with std::sort on the magnitude alone (reference src/filter.cpp:188); the order of exact ties
is part of the result (flood priority, order of the output segments).  The oracle calls the same std::sort on the same list."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from librectify_b200 import api
    e = api.Engine(0, 1)
    yield e
    e.close()


@pytest.fixture(scope="module")
def orc():
    from oracle.bind import Oracle
    return Oracle()


def _cases():
    rng = np.random.default_rng(5)
    for n in (1, 2, 16, 17, 18, 33, 100, 1000, 4097, 19614, 60000):
        for levels in (3, 50, max(2, n // 2), 10 ** 7):
            yield n, (rng.integers(0, levels, n) * 0.37).astype(np.float32), -1
    for n in (40, 700, 5000):                                # the heap-sort branch (depth limit exhausted)
        for d in (0, 1, 3):
            yield n, (rng.integers(0, 9, n) * 1.5).astype(np.float32), d
    yield 3000, np.arange(3000, dtype=np.float32), -1         # sorted the wrong way round, no ties
    yield 3000, np.zeros(3000, np.float32), -1                # one tie group


def test_sort_replay_equals_std_sort(eng, orc):
    steps_total = 0
    for n, vals, depth in _cases():
        want = orc.std_sort_order(vals, depth)
        got, steps = eng.sort_replay(vals, depth)
        steps_total += steps
        assert np.array_equal(got, want), (n, depth, int(np.argmax(got != want)))
    assert steps_total > 0
