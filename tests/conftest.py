import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


@pytest.fixture(scope="session")
def oracle():
    """The CPU restatement (test infrastructure).  Built on demand with gcc."""
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def cfd():
    """The product package; the CUDA library must already be built (python __graft_entry__.py)."""
    import cfd_b200
    if not os.path.exists(cfd_b200.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    cfd_b200.load()
    return cfd_b200
