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


@pytest.fixture(autouse=True)
def bounds_check_report(request):
    """With the bounds-checked build of the library (PTB200_LIB=.../libptb200_chk.so, tools/run_checked.sh) every GPU test ends
    with a look at the device-side failure counters: an index outside its array anywhere in the test fails it.  compute-sanitizer
    is not available on the GPU pool; this is what stands in for memcheck.  No-op for the default build and for CPU tests."""
    yield
    if request.node.get_closest_marker("gpu") is None or os.environ.get("PTB200_CHK_SHRINK"):
        return
    from cuda_pathtracer_b200 import api
    rep = api.check_report()
    if rep is not None:
        assert rep[0] == 0, "bounds-checked build: %d out-of-range indices, first: code %d index %d limit %d" % rep
