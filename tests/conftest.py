import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the reference's host code compiled in the build container)")


@pytest.fixture(scope="session")
def built():
    """Make sure the product library and the CPU oracle are built (both are compiled in-tree)."""
    import __graft_entry__ as g
    g.build(quiet=True)
    return True
