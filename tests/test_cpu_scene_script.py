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
