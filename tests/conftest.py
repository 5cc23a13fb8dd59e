import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    def load(name):
        path = os.path.join(GOLDEN, name + ".npz")
        if not os.path.exists(path):
            pytest.skip(f"golden fixture {name}.npz not generated")
        return np.load(path)
    return load


@pytest.fixture(scope="session")
def pkg():
    import lms_b200
    return lms_b200


@pytest.fixture(scope="session")
def lib(pkg):
    return pkg._lib.load()
