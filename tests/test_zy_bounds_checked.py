"""-m gpu, bounds-checked build only (PTB200_LIB=cuda_pathtracer_b200/libptb200_chk.so; tools/run_checked.sh): the checks themselves.
The default build skips this file.  compute-sanitizer is closed on the GPU pool, so the kernels' indices are verified by a build
that compares every one of them with its array's extent (csrc/pt_device.cuh PT_IDX); tests/conftest.py reads the failure counters
after every GPU test.  Here: a clean run reports nothing, and a limit shrunk on purpose IS reported (a check that cannot fire proves
nothing)."""
import os

import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, api
from util import load_golden_scene

pytestmark = pytest.mark.gpu


def needs_checked_build():
    if not api.bounds_checked_build():
        pytest.skip("default build: run tools/run_checked.sh for the bounds-checked library")


def frames(pt, cam, n):
    pt.HNEE, pt.HCACHE = True, True
    pt.Render(cam, shouldClear=True)
    for _ in range(n):
        pt.Render(cam)
    pt.Finish()


def test_clean_run_reports_nothing():
    needs_checked_build()
    a = load_golden_scene("outside")
    api.check_report()
    pt = api.Pathtracer(a, abi.default_config(160, 120, max_ray_depth=6, interactive_depth=1)).Init()
    frames(pt, a["camera_160x120"], 6)
    hits = pt.debug_trace(a["rays_160x120"], False)
    pt.close()
    assert api.check_report() == (0, 0, 0, 0)
    assert np.array_equal(hits["primitive_id"], a["hits_160x120"]["primitive_id"])      # the checked build computes the same hits


def test_a_shrunk_limit_is_reported():
    needs_checked_build()
    a = load_golden_scene("teapot")
    api.check_report()
    os.environ["PTB200_CHK_SHRINK"] = "1"               # triangle-record extent := 1 for contexts created now
    try:
        pt = api.Pathtracer(a, abi.default_config(160, 120, max_ray_depth=2, interactive_depth=1)).Init()
        pt.debug_trace(a["rays_160x120"], False)
        pt.close()
        n, code, index, limit = api.check_report()
    finally:
        del os.environ["PTB200_CHK_SHRINK"]
    assert n > 0 and code == 3 and index >= 1 and limit == 1          # PT_CHK_WTRI
    assert api.check_report() == (0, 0, 0, 0)                          # reading clears
