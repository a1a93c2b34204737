import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def m4():
    import m4ri_rust_b200
    m4ri_rust_b200.lib()
    return m4ri_rust_b200


@pytest.fixture(scope="session")
def orc():
    import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def gpu(m4):
    if m4.device_count() <= 0:
        pytest.fail("a test marked gpu is running on a box without a usable CUDA device")
    return m4.device_info()
