"""Scene scripts (the reference's getScriptedScene binding, sceneBuilder.h:263-311) and the camera state file
(stateLoader.h:30-84) on the Python mirror.  The scripts below are written for this test in the dialect of the
reference's example_scene.chai / sponza.chai; the expected scene is built by calling the Scene API directly."""
import os

import numpy as np
import pytest

from cuda_pathtracer_b200 import scenes
from cuda_pathtracer_b200.scene import Scene, make_camera, material
from cuda_pathtracer_b200.scene_script import ScriptError, initial_camera, load_scene_script, read_state, run_scene_script, save_state
from util import write_obj

SCRIPT = """
// two cubes: a glass one and a small emitter inside it
var pi = 3.14159
var mat = DiffuseMaterial(make_float3(1, 1, 1))

mat.refractive_index = 1.5
mat.absorption = make_float3(0.1, 0.05, 0.0)
mat.glossy = 0
mat.transmit = 1.0f

var mat_id = scene_add_material(mat)
var model = scene_add_model("box.obj",
            5,
            make_float3(0, 0, 0),
            make_float3(0, 0, 0),
            mat_id, false)
var obj = GameObject(model);
obj.position = make_float3(0, 10, 0);
obj.rotation.y = pi/2
scene_add_object(obj);

var mat2 = DiffuseMaterial(make_float3(1, 1, 1))
mat2.reflect = 0
mat2.emission = make_float3(1.0, 1.0, 1.0)
var mat2_id = scene_add_material(mat2)
var model2 = scene_add_model("box.obj", 5, make_float3(0, 0, 0), make_float3(0, 0, 0), mat2_id, false)

var obj2 = GameObject(model2);
obj2.position = make_float3(0, 10, 0);
obj2.scale = make_float3(0.2, 0.2, 0.2);
obj2.rotation.y = pi/2
obj2.position.x = 1/2 + 0.25 * 2      /* integer division: 0 + 0.5 */
scene_add_object(obj2);
scene_add_plane(Plane(make_float3(0, -1, 0), -3, mat_id))
"""


def _expected(obj_path):
    s = Scene()
    m1 = s.addMaterial(material((1, 1, 1), refractive_index=1.5, absorption=(0.1, 0.05, 0.0), glossy=0, transmit=1.0))
    a = s.addModel(obj_path, 5.0, (0, 0, 0), (0, 0, 0), m1, False)
    s.addObject(a, position=(0, 10, 0), rotation=(0, 3.14159 / 2, 0))
    m2 = s.addMaterial(material((1, 1, 1), reflect=0, emission=(1, 1, 1)))
    b = s.addModel(obj_path, 5.0, (0, 0, 0), (0, 0, 0), m2, False)
    s.addObject(b, position=(0.5, 10, 0), rotation=(0, 3.14159 / 2, 0), scale=(0.2, 0.2, 0.2))
    s.addPlane((0, -1, 0), -3, m1)
    return s.finalize()


def _same(a, b):
    for key in ("vertices", "vertexData", "materials", "instances", "topBvh", "planes", "models"):
        for f in a[key].dtype.names:
            if f == "bvh":
                continue                                  # a host pointer
            assert a[key][f].tobytes() == b[key][f].tobytes(), (key, f)
    assert len(a["bvh"]) == len(b["bvh"]) and all(x.tobytes() == y.tobytes() for x, y in zip(a["bvh"], b["bvh"]))


def test_script_builds_the_same_scene_as_direct_calls(tmp_path):
    write_obj(str(tmp_path / "box.obj"), scenes.cube_soup())
    (tmp_path / "scene.chai").write_text(SCRIPT)
    got = load_scene_script(str(tmp_path / "scene.chai"))
    _same(got.arrays(), _expected(str(tmp_path / "box.obj")).arrays())
    a = got.arrays()
    assert len(a["instances"]) == 2 and len(a["topBvh"]) == 3 and len(a["planes"]) == 1 and len(a["vertices"]) == 24


def test_unknown_names_and_bad_fields_are_errors(tmp_path):
    write_obj(str(tmp_path / "box.obj"), scenes.cube_soup())
    with pytest.raises(ScriptError, match="unknown function 'cene_add_object'"):      # the typo of a half-edited script must not pass silently
        run_scene_script('var m = scene_add_material(DiffuseMaterial(make_float3(1,1,1)))\ncene_add_object(GameObject(0));', str(tmp_path))
    with pytest.raises(ScriptError, match="no field"):
        run_scene_script("var m = DiffuseMaterial(make_float3(1,1,1))\nm.shininess = 3", str(tmp_path))
    with pytest.raises(ScriptError, match="unknown name"):
        run_scene_script("var o = GameObject(model)", str(tmp_path))
    with pytest.raises(ScriptError):
        run_scene_script("var v = make_float3(1, 2)", str(tmp_path))


def test_camera_state_file_round_trip(tmp_path):
    p = str(tmp_path / "save.txt")
    with open(p, "w") as f:
        f.write("9.5|13.25|-1.125\n-0.6|-0.5|0.62449980\n1.5\n8.5\n6.5e-10\n")          # the five lines saveState writes
    cam = read_state(p, 640, 480)
    want = make_camera((9.5, 13.25, -1.125), (-0.6, -0.5, 0.62449980), 1.5, 8.5, 6.5e-10, 640, 480)
    assert cam.tobytes() == want.tobytes()
    save_state(p, cam)
    assert open(p).read().splitlines()[0] == "9.5|13.25|-1.125" and open(p).read().splitlines()[2:] == ["1.5", "8.5", "6.5e-10"]
    assert read_state(p, 640, 480)["eye"].tobytes() == cam["eye"].tobytes()
    assert read_state(str(tmp_path / "missing.txt"), 640, 480).tobytes() == initial_camera(640, 480).tobytes()
    c0 = initial_camera(640, 480)[0]
    assert tuple(c0["eye"]) == (0.0, 2.0, -3.0) and float(c0["focalLength"]) == 5.0


# ---- the reference's OWN scene files through the reference's OWN binding -----------------------------------------------------
# oracle/_ref/libptref_chai.so = getScriptedScene (sceneBuilder.h:271-311) + ChaiScript 6.1.0 as vendored under
# /root/reference/chaiscript, compiled unmodified (oracle/ref_chai.cu).  The scripts are the ones the reference ships.
import ctypes as C
import shutil

from cuda_pathtracer_b200 import abi
from oracle import oracle_py

REF_ROOT = "/root/reference"
CHAI_SO = os.path.join(os.path.dirname(os.path.abspath(oracle_py.__file__)), "_ref", "libptref_chai.so")
needs_ref_chai = pytest.mark.skipif(not (os.path.exists(CHAI_SO) and os.path.exists(os.path.join(REF_ROOT, "example_scene.chai"))),
                                    reason="oracle/_ref/libptref_chai.so or /root/reference not present (build container only)")


class _ChaiRef(oracle_py.RefScene):
    """A RefScene whose Scene was filled by the reference's getScriptedScene."""

    def __init__(self, script_path, cwd):
        self.L = C.CDLL(CHAI_SO, mode=os.RTLD_LOCAL)
        self.L.ref_script_scene.restype = C.c_void_p
        self.L.ref_ptr.restype = C.c_void_p
        old = os.getcwd()
        os.chdir(cwd)                                    # scripts name their models relative to the working directory
        try:
            h = self.L.ref_script_scene(script_path.encode())
        finally:
            os.chdir(old)
        assert h, "the reference's ChaiScript binding rejected " + script_path
        self.h = C.c_void_p(h)


def _compare_scene_arrays(ours, ref):
    for k in ("vertices", "vertexData", "models", "instances", "topBvh", "materials", "spheres", "planes"):
        a, b = ours[k], ref[k]
        assert len(a) == len(b), k
        for f in a.dtype.names:
            if (k, f) in (("models", "bvh"), ("materials", "texture"), ("materials", "normal_texture"), ("vertexData", "tangent"), ("vertexData", "bitangent")):
                continue                                 # host pointers / texture handles; tangents are left uninitialised by the reference without a normal map (scene.h:308)
            assert np.array_equal(np.ascontiguousarray(a[f]).view(np.uint8), np.ascontiguousarray(b[f]).view(np.uint8)), "%s.%s differs" % (k, f)
    assert len(ours["bvh"]) == len(ref["bvh"])
    for x, y in zip(ours["bvh"], ref["bvh"]):
        assert x.tobytes() == y.tobytes()


@pytest.mark.ref
@needs_ref_chai
def test_reference_example_scene_chai_runs_identically(tmp_path):
    """/root/reference/example_scene.chai, with its one typo fixed (line 18 calls `cene_add_object`, which ChaiScript rejects as
    an unknown function -- the shipped file cannot run in the reference either), through the reference's binding and through
    cuda_pathtracer_b200.scene_script: every array Pathtracer::Init would upload must agree byte for byte."""
    text = open(os.path.join(REF_ROOT, "example_scene.chai")).read()
    assert "\ncene_add_object(obj);" in text
    with pytest.raises(ScriptError):                     # unfixed: unknown name, as in ChaiScript
        run_scene_script(text, REF_ROOT)
    text = text.replace("\ncene_add_object(obj);", "\nscene_add_object(obj);")
    shutil.copy(os.path.join(REF_ROOT, "cube.obj"), tmp_path / "cube.obj")
    (tmp_path / "example_scene.chai").write_text(text)
    ref = _ChaiRef("example_scene.chai", str(tmp_path)).arrays()
    ours = load_scene_script(str(tmp_path / "example_scene.chai")).arrays()
    assert len(ref["instances"]) == 2 and len(ref["materials"]) == 2 and len(ref["vertices"]) == 24
    _compare_scene_arrays(ours, ref)


@pytest.mark.ref
@needs_ref_chai
def test_reference_sponza_chai_runs_identically(tmp_path):
    """/root/reference/sponza.chai unmodified.  sponza.obj is not in the mount: a small proxy OBJ (with a texture-free MTL, since
    loadTexture needs a CUDA device, use_cuda.h:108) stands in under the same name; cube.obj is the reference's own."""
    shutil.copy(os.path.join(REF_ROOT, "sponza.chai"), tmp_path / "sponza.chai")
    shutil.copy(os.path.join(REF_ROOT, "cube.obj"), tmp_path / "cube.obj")
    soup = scenes.statue_soup(4, 12) * np.float32(50.0)
    with open(tmp_path / "sponza.obj", "w") as f:
        f.write("mtllib sponza.mtl\n")
        for t in soup.reshape(-1, 3):
            f.write("v %r %r %r\n" % (float(t[0]), float(t[1]), float(t[2])))
        half = len(soup) // 2
        f.write("usemtl stone\n" + "".join("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3) for i in range(half)))
        f.write("usemtl glassy\n" + "".join("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3) for i in range(half, len(soup))))
    (tmp_path / "sponza.mtl").write_text("newmtl stone\nKd 0.7 0.6 0.5\nKs 0.1 0.1 0.1\nd 1\nNs 40\nNi 1\n"
                                         "newmtl glassy\nKd 0.9 0.9 0.9\nKs 0.3 0.3 0.3\nd 0.2\nNs 400\nNi 1.45\n")
    ref = _ChaiRef("sponza.chai", str(tmp_path)).arrays()
    ours = load_scene_script(str(tmp_path / "sponza.chai")).arrays()
    assert len(ref["instances"]) == 2 and len(ref["materials"]) == 4          # white, stone, glassy, light
    _compare_scene_arrays(ours, ref)
