"""CPU tests (no GPU): the C ABI loads and exports what include/*.h declares; POD layouts match the
reference; the CPU oracle reproduces the committed golden vectors that were generated from the
reference's own host code (tests/golden/make_golden.py)."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, api
from oracle import oracle_py
from oracle.oracle_py import OracleLib, oracle_camera
from util import GOLDEN, load_golden_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built):
    L = api.load_library()
    names = set()
    for hdr in ("ptb200.h", "ptb200_scene.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        names |= set(re.findall(r"PT_API\s+[\w\s\*]+?\b(pts?_\w+)\s*\(", src))
    assert len(names) > 40
    missing = [n for n in sorted(names) if not hasattr(L, n)]
    assert not missing, "declared in include/*.h but not exported: %s" % missing


def test_no_device_fails_loudly(built):
    """There is no CPU fallback: without a CUDA device pt_create returns PT_ERR_NO_DEVICE."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = api.load_library()
    h = C.c_void_p()
    cfg = abi.default_config(64, 64)
    assert L.pt_create(C.byref(h), C.byref(cfg)) == -3
    assert b"no CPU fallback" in L.pt_last_error(None)
    with pytest.raises(api.PtError):
        api.Pathtracer(load_golden_scene("teapot"), cfg).Init()


def test_invalid_config_is_rejected(built):
    L = api.load_library()
    h = C.c_void_p()
    for kw in (dict(width=0), dict(world_size=0), dict(rank=2, world_size=2), dict(max_ray_depth=65)):
        base = dict(width=64, height=64)
        base.update(kw)
        cfg = abi.default_config(**base)
        assert L.pt_create(C.byref(h), C.byref(cfg)) == -1


def test_pod_sizes_match_reference_sizeof():
    # sizeof() measured by compiling the reference's types.h; re-measured live when oracle/_ref exists
    expect = {"MATERIAL": 88, "TRIANGLE_V": 48, "TRIANGLE_D": 80, "RADIANCE_CACHE": 112, "BVH_NODE": 32, "MODEL": 32, "INSTANCE": 128,
              "TLAS_NODE": 32, "SPHERE": 20, "PLANE": 20, "RAY": 32, "HIT": 16, "SAMPLE_CACHE": 32, "CAMERA": 76, "TRIANGLE_LIGHT": 8}
    for k, v in expect.items():
        assert getattr(abi, k).itemsize == v, k
    if oracle_py.ref_available(160, 120):
        s = oracle_py.RefLib(160, 120).sizes()
        order = ["MATERIAL", "TRIANGLE_V", "TRIANGLE_D", "RADIANCE_CACHE", "BVH_NODE", "MODEL", "INSTANCE", "TLAS_NODE", "SPHERE", "PLANE",
                 "RAY", "HIT", "SAMPLE_CACHE", "CAMERA"]
        for i, k in enumerate(order):
            assert s[i] == getattr(abi, k).itemsize, k
        assert s[15] == 96      # SceneBuffers (types.h:689-703)


def _oracle_for(a, w, h, **kw):
    cfg = abi.default_config(w, h, **kw)
    o = OracleLib(cfg)
    desc, keep = abi.make_scene_desc(a)
    o.upload_scene(desc, keep)
    return o


@pytest.mark.parametrize("name", ["teapot", "outside"])
def test_oracle_reproduces_reference_vectors(name):
    a = load_golden_scene(name)
    o = _oracle_for(a, 160, 120)
    ci = a["cam_in"]
    cam = oracle_camera(ci[0:3], ci[3:6], ci[6], ci[7], ci[8], 160, 120)
    for f in ("eye", "viewDir", "lt", "u", "v", "d", "focalLength", "aperture"):
        assert np.array_equal(np.asarray(cam[0][f]).view(np.uint32), np.asarray(a["camera_160x120"][0][f]).view(np.uint32)), f
    rays = o.primary_rays(cam, 0)
    assert rays.tobytes() == a["rays_160x120"].tobytes()
    assert o.trace(rays, False).tobytes() == a["hits_160x120"].tobytes()
    assert o.trace(rays, True).tobytes() == a["anyhits_160x120"].tobytes()
    o.close()


def test_oracle_reproduces_teapot_known_answers():
    """BASELINE.md section 4 / SURVEY.md 8(c): 640x480, camera (0,2,-6)->(0,0,1), d 1.5, focal 5, aperture 0.01."""
    ka = json.load(open(os.path.join(GOLDEN, "known_answers.json")))
    a = load_golden_scene("teapot")
    assert len(a["vertices"]) == ka["triangles"] == 6320 and len(a["bvh"][0]) == ka["bvh_nodes"] == 4129
    leaves = a["bvh"][0]["count"] > 0
    assert leaves.sum() == 2065 and a["bvh"][0]["count"].max() == 4
    o = _oracle_for(a, 640, 480)
    cam = oracle_camera((0, 2, -6), (0, 0, 1), 1.5, 5, 0.01, 640, 480)
    hits = o.trace(o.primary_rays(cam, 0), False)
    hit = hits["primitive_id"] != 0xFFFFFFFF
    assert int(hit.sum()) == ka["hit_pixels"] == 41550
    assert int(hits["primitive_id"][hit].astype(np.uint64).sum()) == ka["sum_primitive_id"] == 84689855
    assert abs(float(hits["t"][hit].astype(np.float64).sum()) - ka["sum_t"]) < 1e-6
    o.close()


def test_oracle_frame_loop_bookkeeping():
    """Render(shouldClear) semantics, pathtracer.h:239-301: interactive_depth samples on a clearing frame, one per
    converge frame, randIdx bumped once per primary launch and once per bounce, image.w = sample count."""
    a = load_golden_scene("teapot")
    o = _oracle_for(a, 64, 48, max_ray_depth=3, interactive_depth=2)
    cam = oracle_camera((0, 2, -6), (0, 0, 1), 1.5, 5, 0.01, 64, 48)
    o.render(cam, True)
    s = C.c_uint32()
    r = C.c_uint32()
    o.L.ora_get_rand_state(o.h, C.byref(s), C.byref(r))
    assert (s.value, r.value) == (2, 2 * (1 + 2))        # HNEE true -> max_bounces = interactive_depth
    o.render(cam, False)
    o.L.ora_get_rand_state(o.h, C.byref(s), C.byref(r))
    assert (s.value, r.value) == (3, 6 + 1 + 3)
    img = o.read_image()
    assert np.all(img[..., 3] == 3.0) and np.all(img[..., :3] >= 0) and np.isfinite(img).all()
    o.set_toggles(False, True)
    o.render(cam, True)                                   # NEE off -> interactive_depth + 1 bounces
    o.L.ora_get_rand_state(o.h, C.byref(s), C.byref(r))
    assert (s.value, r.value) == (2, 2 * (1 + 3))
    o.close()
