"""Device BVH2 build (pt_build_bvh) against the host builder, which the CPU suite pins against the
reference's createBVHBinned (tests/test_cpu_host_scene.py): same nodes, same triangle order, same node numbering.
Integer fields are compared bit for bit; box floats by value (a zero may differ in sign, see bvh_build.cu)."""
import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, api, scenes
from cuda_pathtracer_b200.scene import Scene, material

pytestmark = pytest.mark.gpu


def host_and_device(verts):
    out = []
    for dev in (-1, 0):
        s = Scene().set_build_device(dev)
        s.addMaterial(material())
        s.addModelSoup(verts)
        s.addObject(0)
        s.finalize()
        out.append((s.arrays(), s.last_build_ms()))
    return out


def assert_same_tree(h, d):
    hb, db = h["bvh"][0], d["bvh"][0]
    assert len(hb) == len(db), (len(hb), len(db))
    for f in ("first", "count"):
        bad = np.nonzero(hb[f] != db[f])[0]
        assert len(bad) == 0, "%s differs at node %d of %d: host %s device %s" % (f, bad[0], len(hb), hb[bad[0]], db[bad[0]])
    for f in ("vmin", "vmax"):
        bad = np.nonzero((hb[f] != db[f]).any(axis=1))[0]
        assert len(bad) == 0, "%s differs at node %d: host %s device %s" % (f, bad[0], hb[bad[0]], db[bad[0]])
    for f in ("v0", "v1", "v2"):
        bad = np.nonzero((h["vertices"][f] != d["vertices"][f]).any(axis=1))[0]
        assert len(bad) == 0, "triangle order differs first at position %d of %d" % (bad[0], len(h["vertices"]))
    for f in h["vertexData"].dtype.names:                       # field by field: the structs carry padding bytes
        assert h["vertexData"][f].tobytes() == d["vertexData"][f].tobytes(), f
    assert np.array_equal(h["models"]["nrBvhNodes"], d["models"]["nrBvhNodes"])
    for f in h["topBvh"].dtype.names:
        assert np.array_equal(h["topBvh"][f], d["topBvh"][f]), f


def _random_soup(n, spread, size, seed=11):
    rng = np.random.default_rng(seed)
    return (rng.random((n, 1, 3), np.float32) * spread + (rng.random((n, 3, 3), np.float32) - 0.5) * size).reshape(-1, 9)


def _coplanar_grid():                      # coplanar, many SAH ties, zero coordinates
    g = np.stack(np.meshgrid(np.arange(64, dtype=np.float32), np.arange(64, dtype=np.float32), indexing="ij"), -1).reshape(-1, 2)
    flat = np.zeros((len(g), 9), np.float32)
    flat[:, 0], flat[:, 2] = g[:, 0], g[:, 1]
    flat[:, 3], flat[:, 5] = g[:, 0] + 1, g[:, 1]
    flat[:, 6], flat[:, 8] = g[:, 0], g[:, 1] + 1
    return flat


SOUPS = {
    "random-soup-20k": lambda: _random_soup(20000, 10, 1.0),
    "cathedral": scenes.cathedral_soup,
    "statue": scenes.statue_soup,
    "cube": scenes.cube_soup,
    "heightfield-200x201": lambda: scenes.heightfield_soup(200, 201),
    "voxel-terrain-64": lambda: scenes.voxel_terrain_soup(64)[0],
    "identical-x40": lambda: np.tile(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), (40, 1)),     # centroid extent 0: one big leaf
    "coplanar-grid-4096": _coplanar_grid,
    "random-soup-300k": lambda: _random_soup(300000, 50, 0.3),
}
for _n in (1, 2, 4, 5, 6, 9, 17):
    SOUPS["tiny-%d" % _n] = (lambda n: (lambda: _random_soup(n, 4, 1.0, seed=n)))(_n)


@pytest.mark.parametrize("name", list(SOUPS))
def test_device_build_equals_host_build(name):
    verts = np.ascontiguousarray(SOUPS[name](), np.float32).reshape(-1, 9)
    (h, hms), (d, dms) = host_and_device(verts)
    assert_same_tree(h, d)


def test_raw_entry_point_order_and_offsets():
    """pt_build_bvh itself: `order` is a permutation, leaves tile [offset, offset+n) in order, inner nodes point at adjacent
    children handed out in pre-order (bvhBuilder.h:222-223)."""
    v = scenes.statue_soup().reshape(-1, 9)
    nodes, order, ms, levels = api.build_bvh(v, tri_offset=1000)
    n = len(v)
    assert sorted(order.tolist()) == list(range(n)) and levels > 3 and ms > 0
    leaves = nodes[nodes["count"] > 0]
    assert int(leaves["count"].sum()) == n and int(leaves["first"].min()) == 1000 and int((leaves["first"] + leaves["count"]).max()) == 1000 + n
    inner = nodes[nodes["count"] == 0]
    assert len(nodes) == 2 * len(inner) + 1
    assert np.array_equal(np.sort(inner["first"]), 1 + 2 * np.arange(len(inner), dtype=np.uint32))
    # every triangle of a leaf lies inside the leaf's box
    t = np.zeros(n, abi.TRIANGLE_V); t["v0"], t["v1"], t["v2"] = v[:, 0:3], v[:, 3:6], v[:, 6:9]
    t = t[order]
    for leaf in leaves[:: max(1, len(leaves) // 200)]:
        tri = t[leaf["first"] - 1000: leaf["first"] - 1000 + leaf["count"]]
        pts = np.concatenate([tri["v0"], tri["v1"], tri["v2"]])
        assert (pts >= leaf["vmin"]).all() and (pts <= leaf["vmax"]).all()


def test_render_with_device_built_scene_is_bit_identical():
    """The whole path on a device-built tree: same image as with the host-built tree."""
    imgs = []
    for dev in (-1, 0):
        s = Scene().set_build_device(dev)
        s.addMaterial(material(diffuse=(0.7, 0.6, 0.5)))
        s.addModelSoup(scenes.statue_soup())
        s.addObject(0)
        s.finalize()
        cfg = abi.default_config(160, 120, max_ray_depth=4, interactive_depth=1)
        pt = api.Pathtracer(s.arrays(), cfg).Init()
        cam = scenes.camera_sweep("sibenik", 160, 120, 4)[0]
        pt.Render(cam, shouldClear=True)
        for _ in range(4):
            pt.Render(cam)
        pt.Finish()
        imgs.append(pt.read_image())
        pt.close()
    assert imgs[0].tobytes() == imgs[1].tobytes()


WIDE_CASES = ["cube", "statue", "cathedral", "identical-x40", "coplanar-grid-4096", "tiny-1", "tiny-4", "tiny-5", "tiny-17", "heightfield-200x201", "random-soup-300k"]


@pytest.mark.parametrize("name", WIDE_CASES)
def test_device_wide_bvh_equals_host_converter(name):
    """BVH2 -> 8-wide conversion on the device (default of pt_upload_scene) against the host converter
    (PT_FLAG_HOST_WIDE_BVH): byte-identical node and triangle-record arrays."""
    verts = np.ascontiguousarray(SOUPS[name](), np.float32).reshape(-1, 9)
    s = Scene()
    s.addMaterial(material())
    s.addModelSoup(verts)
    s.addModelSoup(scenes.cube_soup())                      # a second model: node / triangle bases are absolute
    s.addObject(0)
    s.addObject(1, position=(3, 0, 0))
    s.finalize()
    out = []
    for flags in (abi.FLAG_HOST_WIDE_BVH, 0):
        pt = api.Pathtracer(s.arrays(), abi.default_config(64, 64, max_ray_depth=2, interactive_depth=1, flags=flags)).Init()
        out.append(pt.debug_read_wide_bvh())
        pt.close()
    (hn, ht), (dn, dt) = out
    assert hn.shape == dn.shape and ht.shape == dt.shape, (hn.shape, dn.shape, ht.shape, dt.shape)
    bad = np.nonzero((hn != dn).any(axis=1))[0]
    assert len(bad) == 0, "wide node %d of %d differs: host %s device %s" % (bad[0], len(hn), hn[bad[0]], dn[bad[0]])
    bad = np.nonzero((ht != dt).any(axis=1))[0]
    assert len(bad) == 0, "triangle record %d of %d differs" % (bad[0], len(ht))


def _instanced_scene(n, seed):
    rng = np.random.default_rng(seed)
    s = Scene()
    s.addMaterial(material())
    s.addModelSoup(scenes.cube_soup())
    s.addModelSoup(scenes.statue_soup()[:2000])
    for i in range(n):
        s.addObject(int(i % 2), position=rng.uniform(-20, 20, 3), rotation=rng.uniform(-3, 3, 3), scale=rng.uniform(0.3, 2.5, 3))
    s.finalize()
    return s, rng


@pytest.mark.parametrize("n", [1, 2, 3, 10, 33, 200])
def test_device_top_level_bvh_equals_host_build(n):
    """k_build_tlas against the host BuildTopLevelBVH (pinned to the reference in tests/test_cpu_host_scene.py): the same
    nodes bit for bit, also after the objects moved (Scene::update, scene.h:367-401)."""
    s, rng = _instanced_scene(n, 100 + n)
    a = s.arrays()
    pt = api.Pathtracer(a, abi.default_config(64, 64, max_ray_depth=2, interactive_depth=1)).Init()
    for step in range(3):
        pt.update_instances_device_tlas(a["instances"])
        got, want = pt.debug_read_tlas(), a["topBvh"]
        assert len(got) == 2 * n - 1
        for f in ("vmin", "vmax", "children"):
            bad = np.nonzero((got[f] != want[f]).reshape(len(got), -1).any(axis=1))[0]
            assert len(bad) == 0, "step %d: %s differs at node %d: device %s host %s" % (step, f, bad[0], got[bad[0]], want[bad[0]])
        leaves = want["children"] == 0
        assert np.array_equal(got["leaf"][leaves], want["leaf"][leaves])
        for i in range(n):                                   # move everything, rebuild on the host, compare again
            s.setObject(i, rng.uniform(-20, 20, 3), rng.uniform(-3, 3, 3), rng.uniform(0.3, 2.5, 3))
        s.update()
        a = s.arrays()
    pt.close()


def test_render_with_device_top_level_bvh_is_bit_identical():
    imgs = []
    for device_tlas in (False, True):
        sc = scenes.outside_scene()
        pt = api.Pathtracer(sc.arrays(), abi.default_config(160, 120, max_ray_depth=6, interactive_depth=1)).Init()
        pt.DEVICE_TLAS = device_tlas
        cam = scenes.camera_sweep("outside", 160, 120, 4)[1]
        pt.Render(cam, shouldClear=True)
        for _ in range(4):
            pt.Render(cam)
        pt.Finish()
        imgs.append(pt.read_image())
        pt.close()
    assert imgs[0].tobytes() == imgs[1].tobytes() and float(imgs[0][..., :3].sum()) > 0


def test_scene_script_renders_headless(tmp_path):
    """examples/render_script.py: scene script + camera state file -> PFM (the reference's `pathtracer scene.chai` flow)."""
    import subprocess, sys, os
    from util import read_pfm, write_obj
    write_obj(str(tmp_path / "box.obj"), scenes.cube_soup())
    (tmp_path / "scene.chai").write_text('var m = DiffuseMaterial(make_float3(0.8, 0.7, 0.6))\nvar id = scene_add_material(m)\n'
                                         'var box = scene_add_model("box.obj", 2, make_float3(0,0,0), make_float3(0,0,0), id, false)\n'
                                         'var o = GameObject(box)\no.position = make_float3(0, 2, 4)\no.rotation.y = 0.5\nscene_add_object(o);\n')
    root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "render_script.py"), str(tmp_path / "scene.chai"), str(tmp_path / "o.pfm"),
                        "--state", str(tmp_path / "none.txt"), "--spp", "8", "--size", "160x120", "--depth", "4"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-1500:]
    img = read_pfm(str(tmp_path / "o.pfm"))
    assert img.shape[:2] == (120, 160) and np.isfinite(img).all() and img.std() > 0.01          # the box is in view against the white sky
