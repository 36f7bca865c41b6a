import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU restatement (checker). Built on demand from oracle/rectify_oracle.cpp."""
    from oracle.bind import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref_canon():
    """The compiled, unmodified reference (strict-IEEE, fixed seed). Only where oracle/_ref was built."""
    from oracle.bind import Ref, ref_available
    if not ref_available("canon"):
        pytest.skip("oracle/_ref/librectify_canon.so not built (needs /root/reference; make -C oracle ref)")
    return Ref("canon")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    d = os.path.join(ROOT, "tests", "golden")

    def load(name):
        return dict(np.load(os.path.join(d, name + ".npz"), allow_pickle=False))
    return load
