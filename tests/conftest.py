import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle as po

    return po.oracle()


@pytest.fixture(scope="session")
def engine():
    """The product engine on cuda:0.  Fails (does not skip) when the extension or the GPU is missing."""
    import bitcoincashplus_b200 as bcp

    return bcp.Engine(0)
