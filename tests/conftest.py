import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    from oracle import loader
    if not loader.available("port"):
        loader.build("port")
    return loader.load("port")


@pytest.fixture(scope="session")
def ref():
    from oracle import loader
    if not loader.available("ref"):
        if os.path.isdir("/root/reference/include/cstone"):
            loader.build("ref")
    if not loader.available("ref"):
        pytest.skip("oracle/_ref/liboracle_ref.so not built (reference tree absent)")
    return loader.load("ref")


@pytest.fixture(scope="session")
def oracle():
    """the strongest CPU checker present: the reference build if it travelled, else our restatement"""
    from oracle import loader
    if loader.available("ref"):
        return loader.load("ref")
    if not loader.available("port"):
        loader.build("port")
    return loader.load("port")
