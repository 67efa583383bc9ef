import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    from _libs import oracle
    return oracle()


@pytest.fixture(scope="session")
def reflib():
    from _libs import ref
    lib = ref()
    if lib is None:
        pytest.skip("oracle/_ref/libsrslte_ref.so not available (needs /root/reference to build)")
    return lib
