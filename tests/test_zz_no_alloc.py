"""-m gpu, collected last: pt_trace_static (include/ptb200.h) -- the traversal hook whose scratch is a file-scope device array
instead of cudaMalloc'd buffers -- against the existing entry point pt_debug_trace, through the C ABI.

No counterpart in the reference; the entry point was added for callers that cannot
allocate.  Hit records are integer / index work plus one float t computed by the same kernel with the same arithmetic, so
the bar is the one of tests/test_gpu_trace.py: bit-exact.  Two calls in a row (the second on fewer rays, so it sees the
first call's leftovers in the shared array and a non-zero cursor), nearest-hit and any-hit, and the size limit."""
import ctypes as C

import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, api
from util import load_golden_scene

pytestmark = pytest.mark.gpu


def make_rays(a, w, h, seed):
    """Random rays from inside (and just around) the bounding box of the scene's triangles."""
    rng = np.random.default_rng(seed)
    n = w * h
    rays = np.zeros(n, abi.RAY)
    v = np.concatenate([a["vertices"]["v0"], a["vertices"]["v1"], a["vertices"]["v2"]])
    lo, hi = v.min(0), v.max(0)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    rays["origin"] = rng.uniform(lo - 1.0, hi + 1.0, (n, 3)).astype(np.float32)
    rays["length"] = np.float32(1e30)
    rays["direction"] = d
    rays["pixeli"] = np.arange(n, dtype=np.uint32)
    return rays


def same(a, b):
    hit = a["primitive_id"] != 0xFFFFFFFF
    assert np.array_equal(a["primitive_id"], b["primitive_id"])
    assert np.array_equal(a["primitive_type"][hit], b["primitive_type"][hit]) and np.array_equal(a["instance_id"][hit], b["instance_id"][hit])
    assert np.array_equal(a["t"][hit].view(np.uint32), b["t"][hit].view(np.uint32))
    return int(hit.sum())


def test_static_scratch_trace_matches_debug_trace_twice_in_a_row():
    a = load_golden_scene("teapot")
    cfg = abi.default_config(320, 240, max_ray_depth=2, interactive_depth=1)
    pt = api.Pathtracer(a, cfg).Init()
    try:
        cam_rays = a["rays_160x120"]                          # the golden camera rays of the scene (tests/golden)
        big = np.concatenate([cam_rays, make_rays(a, 320, 240, 1), make_rays(a, 320, 240, 2)])      # 172,800 rays: wavefront form
        small = make_rays(a, 97, 31, 3)                                                           # 3,007 rays: cooperative form, ragged size
        total_hits = 0
        for any_hit in (False, True):
            for rays in (big, small):                       # second call: fewer rays over the first call's leftovers
                want = pt.debug_trace(rays, any_hit)
                got = pt.trace_static(rays, any_hit)
                total_hits += same(got, want)
                again = pt.trace_static(rays, any_hit)      # and the same call twice in a row
                same(again, want)
        assert total_hits > 10000
        # one ray, and the full limit
        one = pt.trace_static(cam_rays[9680:9681], False)
        same(one, pt.debug_trace(cam_rays[9680:9681], False))
        full = np.resize(big, abi.STATIC_TRACE_MAX_RAYS)
        same(pt.trace_static(full, False), pt.debug_trace(full, False))
        # beyond the limit: a clear error, nothing launched
        over = np.resize(big, abi.STATIC_TRACE_MAX_RAYS + 1)
        out = np.zeros(len(over), abi.HIT)
        rc = pt.L.pt_trace_static(pt.h, abi.ptr(over), len(over), 0, abi.ptr(out))
        assert rc == -1
        pt.L.pt_last_error.restype = C.c_char_p
        msg = pt.L.pt_last_error(pt.h).decode()
        assert "PT_STATIC_TRACE_MAX_RAYS" in msg and str(abi.STATIC_TRACE_MAX_RAYS) in msg
        assert not out["primitive_id"].any()
        # the context still works afterwards
        same(pt.trace_static(small, False), pt.debug_trace(small, False))
    finally:
        pt.close()


def test_static_scratch_is_shared_between_contexts():
    """Two contexts (two scenes) on the device use the same array one after the other."""
    a, b = load_golden_scene("teapot"), load_golden_scene("outside")
    cfg = abi.default_config(160, 120, max_ray_depth=2, interactive_depth=1)
    p, q = api.Pathtracer(a, cfg).Init(), api.Pathtracer(b, cfg).Init()
    try:
        ra, rb = make_rays(a, 160, 120, 5), make_rays(b, 200, 90, 6)
        for _ in range(2):
            same(p.trace_static(ra), p.debug_trace(ra))
            same(q.trace_static(rb, True), q.debug_trace(rb, True))
    finally:
        p.close(); q.close()
