import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py

    oracle_py.load()
    return oracle_py


@pytest.fixture(scope="session")
def pcp():
    """One device context for the GPU tests.  No fallback: a missing library or GPU is an error."""
    from pointcloudprocessing_b200 import capi

    ctx = capi.Pcp(0)
    yield ctx
    ctx.close()
