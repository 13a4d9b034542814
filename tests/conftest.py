import importlib
import os
import sys

import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pt():
    return importlib.import_module("hiprt-path-tracer_b200")


@pytest.fixture(scope="session")
def wl():
    return importlib.import_module("hiprt-path-tracer_b200.workloads")


@pytest.fixture(scope="session")
def ref():
    """ORACLE (test infrastructure): the reference's CPU kernels, oracle/_ref."""
    from oracle import ref as r
    if not r.available("mis"):
        pytest.skip("oracle/_ref is not built (python oracle/build_ref.py needs /root/reference)")
    return r


@pytest.fixture(scope="session")
def ctx(pt):
    c = pt.Context(0)
    yield c
    c.close()


GOLDEN = os.path.join(ROOT, "tests", "golden")
