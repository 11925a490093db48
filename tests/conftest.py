import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_sessionstart(session):
    """Built artefacts are git-ignored: in a fresh checkout build them once (nvcc cross-compiles without a GPU)."""
    need = [os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_b200.so"), os.path.join(ROOT, "oracle", "libbcp_oracle.so"),
            os.path.join(ROOT, "tests", "hostsim", "libhostsim.so"), os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_chain.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__

        __graft_entry__.build()


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle as po

    return po.oracle()


@pytest.fixture(scope="session")
def engine():
    """The product engine on cuda:0.  Fails (does not skip) when the extension or the GPU is missing."""
    import bitcoincashplus_b200 as bcp

    return bcp.Engine(0)
