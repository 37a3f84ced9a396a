import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "hydragnn-gps-support_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import torch

    import parity_util

    return parity_util.golden()


@pytest.fixture(scope="session")
def libhgk():
    """Build (if needed) and load the C-ABI library."""
    sys.path.insert(0, ROOT)
    import __graft_entry__

    __graft_entry__.build()
    from hydragnn import _hgk

    return _hgk.lib()
