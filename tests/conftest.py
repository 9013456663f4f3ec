import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_mod():
    """The CPU checker (test infrastructure): builds the C restatement if needed."""
    from oracle import oracle

    oracle.build(("port",))
    if os.path.isdir("/root/reference/lib"):
        oracle.build(("ref", "tskit"))
    return oracle


@pytest.fixture(scope="session")
def native_built():
    from msprime_b200 import build

    build.build()
    return True
