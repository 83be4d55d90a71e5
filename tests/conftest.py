import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    """The CPU checker: the reference's own sources when oracle/_ref is present, else the port."""
    from oracle import oracle as O
    O.build()
    return O.load("auto")


@pytest.fixture(scope="session")
def gpu_lib():
    import simplefe_b200
    from simplefe_b200 import build
    build.build()
    l = simplefe_b200.lib()
    n = l.sfe_device_count()
    assert n > 0, "GPU test on a box without a CUDA device: " + simplefe_b200.last_error()
    return l
