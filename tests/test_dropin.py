"""The drop-in boundary exercised from the reference's side: oracle/_ref/dropin_160x120 is the reference's OWN host
code (Scene::addModel with its tinyobjloader, createBVHBinned, BuildTopLevelBVH, Camera -- compiled from
/root/reference/src where it lies) driving libptb200.so through include/ptb200_pathtracer.h, the
`class Pathtracer : public Application` adapter.  The binary is prebuilt by __graft_entry__.build() in the build
container (oracle/_ref/ travels to the GPU box); the OBJ it loads is generated here."""
import os
import subprocess

import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, scenes
from cuda_pathtracer_b200.scene import Scene, make_camera, material
from util import DROPIN, read_pfm, write_obj

needs_dropin = pytest.mark.skipif(not os.path.exists(DROPIN), reason="oracle/_ref/dropin_160x120 not built (needs /root/reference at build time)")
EYE = (1.0, 3.0, -7.0)


def _mesh(tmp_path):
    tri = scenes.statue_soup(4, 40) - np.tile(np.array([3.0, 0.0, 4.0], np.float32), 3)      # 3200 triangles around the origin
    p = str(tmp_path / "blob.obj")
    write_obj(p, np.ascontiguousarray(tri, np.float32))
    return p


@needs_dropin
def test_dropin_binary_fails_loudly_without_a_device(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([DROPIN, _mesh(tmp_path), str(tmp_path / "o.pfm"), "1"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "no CUDA device" in r.stderr and not os.path.exists(tmp_path / "o.pfm")


@needs_dropin
@pytest.mark.gpu
def test_reference_host_code_renders_through_the_adapter(tmp_path):
    """Reference Scene + adapter + libptb200 (PFM) vs the CPU oracle fed by this repo's own host scene code on the
    same OBJ, camera, seeds, 1 clearing + 6 converge samples, NEE + guiding on (emissive second instance)."""
    from oracle.oracle_py import OracleLib
    obj = _mesh(tmp_path)
    out = str(tmp_path / "dropin.pfm")
    r = subprocess.run([DROPIN, obj, out, "6", str(EYE[0]), str(EYE[1]), str(EYE[2]), "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    img = read_pfm(out)
    assert img.shape == (120, 160, 3) and np.isfinite(img).all()
    # the same scene through this repo's host code (pts_* C++), rendered by the oracle
    s = Scene()
    white = s.addMaterial(material((0.7, 0.6, 0.5)))
    lamp = s.addMaterial(material((1, 1, 1), emission=(10, 10, 10)))
    m = s.addModel(obj, 1.0, (0, 0, 0), (0, 0, 0), white, False)
    s.addObject(m)
    s.addObject(m, material=lamp, position=(0, 6, 0), scale=(0.3, 0.3, 0.3))
    s.addPlane((0, -1, 0), -3, white)
    a = s.finalize().arrays()
    cfg = abi.default_config(160, 120, max_ray_depth=32, interactive_depth=1)          # MAX_RAY_DEPTH of the reference build (constants.h:13)
    d = np.array([0.0, 1.0, 0.0]) - np.array(EYE)
    cam = make_camera(EYE, d / np.linalg.norm(d), 1.5, 5.0, 0.01, 160, 120)
    ora = OracleLib(cfg)
    desc, keep = abi.make_scene_desc(a)
    ora.upload_scene(desc, keep)
    ora.set_toggles(True, True)
    ora.render(cam, True)
    for _ in range(6):
        ora.render(cam, False)
    ref = ora.read_image()
    ref = ref[..., :3] / ref[..., 3:4]
    err = float(np.sqrt(np.mean((img.astype(np.float64) - ref) ** 2)) / np.sqrt(np.mean(ref.astype(np.float64) ** 2)))
    assert err <= 2e-2, err
    ora.close()


EXAMPLE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "examples", "render_pfm")


@needs_dropin
@pytest.mark.gpu
def test_adapter_with_device_top_level_bvh_gives_the_same_picture(tmp_path):
    """Pathtracer::deviceTopLevelBVH: the clearing frame rebuilds the top-level BVH on the device from scene.instances (the
    host tree is scrambled after Init, so a render that still read it would differ)."""
    obj = _mesh(tmp_path)
    outs = []
    for dev in ("0", "1"):
        out = str(tmp_path / ("tlas%s.pfm" % dev))
        r = subprocess.run([DROPIN, obj, out, "4", str(EYE[0]), str(EYE[1]), str(EYE[2]), "1", dev], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1]


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(EXAMPLE), reason="examples/render_pfm not built")
def test_cpp_only_headless_render_to_pfm(tmp_path):
    """examples/render_pfm.cpp: the C ABI from plain C++ (pts_* scene code + pt_* render path), PFM out; the same scene
    through the Python mirror must give the identical picture (guiding off would be bit-exact; it is on by default,
    so float atomics leave ~1e-9)."""
    from cuda_pathtracer_b200 import api
    obj = _mesh(tmp_path)
    out = str(tmp_path / "cpp.pfm")
    r = subprocess.run([EXAMPLE, obj, out, "160", "120", "5", "8", "1", "0"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    img = read_pfm(out)
    s = Scene()
    white = s.addMaterial(material((0.7, 0.7, 0.7)))
    lamp = s.addMaterial(material((1, 1, 1), emission=(12, 12, 12)))
    m = s.addModel(obj, 1.0, (0, 0, 0), (0, 0, 0), white, False)
    s.addObject(m)
    s.addObject(m, material=lamp, position=(0, 7, 0), scale=(0.3, 0.3, 0.3))
    s.addPlane((0, -1, 0), -3, white)
    cfg = abi.default_config(160, 120, max_ray_depth=8, interactive_depth=1)
    pt = api.Pathtracer(s.finalize(), cfg).Init()
    eye = np.array([1.0, 3.0, -8.0]); d = np.array([0.0, 1.0, 0.0]) - eye
    cam = make_camera(tuple(eye), d / np.float32(np.sqrt(np.float32(d @ d))), 1.5, 5.0, 0.01, 160, 120)
    pt.Render(cam, shouldClear=True)
    for _ in range(5):
        pt.Render(cam)
    pt.Finish()
    ref = pt.read_image(); ref = ref[..., :3] / ref[..., 3:4]
    pt.close()
    err = float(np.sqrt(np.mean((img.astype(np.float64) - ref) ** 2)) / np.sqrt(np.mean(ref.astype(np.float64) ** 2)))
    assert err < 1e-3, err
    # and with the display pass
    r = subprocess.run([EXAMPLE, obj, out, "160", "120", "5", "8", "1", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and np.isfinite(read_pfm(out)).mean() > 0.99
