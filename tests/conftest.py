import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as ge
    return ge.load()


@pytest.fixture(scope="session")
def ob():
    from oracle import binding
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        binding.build()
    return binding


@pytest.fixture(scope="session")
def ur5(pkg):
    return pkg.ModelDescription.load("ur5_boxes8.json")


@pytest.fixture(scope="session")
def ur5_oracle(ob, ur5):
    return ob.Oracle(ur5)


@pytest.fixture(scope="session")
def golden_configs():
    """The 220 UR5 configurations of the reference's tests/random-numbers.hh (m1), extracted by
    tests/golden/make_golden.py."""
    return np.load(os.path.join(ROOT, "tests", "golden", "ur5_random_numbers_m1.npy"))


def has_gpu():
    try:
        import __graft_entry__ as ge
        return ge.load().lib().hppb200_device_count() > 0
    except Exception:
        return False
