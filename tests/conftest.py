import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    from oracle.pyoracle import Oracle, build, available
    if not available("orc"):
        build()
    return Oracle("orc")


@pytest.fixture(scope="session")
def ref():
    from oracle.pyoracle import Oracle, available
    if not available("ref"):
        pytest.skip("oracle/_ref/libgravomg_ref.so not built (needs /root/reference)")
    return Oracle("ref")


@pytest.fixture(scope="session")
def ctx():
    """The CUDA path through the C ABI. Fails loudly if the library or the GPU is missing."""
    from gravomg_b200.capi import Context
    c = Context(0)
    yield c
    c.close()
