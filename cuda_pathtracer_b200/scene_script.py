"""Scene scripts and the camera state file.

The reference loads a scene from a ChaiScript file through ``getScriptedScene`` (reference src/sceneBuilder.h:263-311),
which binds a handful of names onto its ``Scene``: ``DiffuseMaterial``, ``make_float3``, ``GameObject``, ``Plane``,
``scene_add_material / scene_add_model / scene_add_plane / scene_add_object`` and the fields of ``Material``,
``GameObject`` and ``float3``.  C++ host code keeps using the reference's own binding (it only fills a ``Scene``; the
adapter in include/ptb200_pathtracer.h sits below it).  This module is the same binding for the Python mirror: an
interpreter for the statement forms those scene files use -- ``var x = expr``, ``a.b.c = expr``, calls, arithmetic,
numbers with an ``f`` suffix, strings, ``true``/``false``, ``//`` comments, optional ``;`` -- evaluated against
``cuda_pathtracer_b200.scene.Scene``.  An unknown name is an error, as it is in ChaiScript.

``read_state`` / ``save_state`` / ``initial_camera`` mirror reference src/stateLoader.h:30-84 (five lines: eye ``x|y|z``,
viewDir ``x|y|z``, d, focalLength, aperture).
"""
import os
import re

import numpy as np

from .scene import Scene, make_camera, material


class ScriptError(Exception):
    pass


class Float3:
    """float3 with assignable x / y / z (sceneBuilder.h:301-303)."""

    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)

    def tuple(self):
        return (self.x, self.y, self.z)

    def copy(self):
        return Float3(self.x, self.y, self.z)


class ScriptMaterial:
    """Material (types.h:33-56) with the fields the binding exposes (sceneBuilder.h:290-297)."""
    FIELDS = ("diffuse_color", "specular_color", "emission", "reflect", "glossy", "transmit", "refractive_index", "absorption")

    def __init__(self, color):                      # Material::DIFFUSE, types.h:48-55
        self.diffuse_color, self.specular_color = color.copy(), Float3()
        self.emission, self.absorption = Float3(), Float3()
        self.reflect = self.glossy = self.transmit = self.refractive_index = 0.0

    def record(self):
        return material(self.diffuse_color.tuple(), self.specular_color.tuple(), self.emission.tuple(), self.reflect, self.glossy,
                        self.transmit, self.refractive_index, self.absorption.tuple())


class ScriptObject:
    """GameObject(model_id, material_id = 0xffffffff), types.h:416-429."""
    FIELDS = ("position", "rotation", "scale")

    def __init__(self, model_id, material_id=0xFFFFFFFF):
        self.model_id, self.material_id = int(model_id), int(material_id)
        self.position, self.rotation, self.scale = Float3(), Float3(), Float3(1, 1, 1)


class ScriptPlane:
    def __init__(self, normal, d, mat):              # Plane(float3, int, int), sceneBuilder.h:282
        self.normal, self.d, self.material = normal.copy(), float(d), int(mat)


_TOKEN = re.compile(r"""\s*(?:(?P<num>(?:\d+\.\d*|\.\d+|\d+)(?:[eE][+-]?\d+)?f?)|(?P<str>"(?:[^"\\]|\\.)*")|(?P<id>[A-Za-z_]\w*)|(?P<op>==|!=|<=|>=|[-+*/().,;={}<>]))""")


def _tokenize(text):
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out, pos = [], 0
    while pos < len(text):
        if text[pos:].strip() == "":
            break
        m = _TOKEN.match(text, pos)
        if not m:
            raise ScriptError("cannot read the script at: %r" % text[pos:pos + 30].strip())
        nl = "\n" in text[pos:m.start(m.lastgroup)]
        out.append((m.lastgroup, m.group(m.lastgroup), nl))
        pos = m.end()
    out.append(("end", "", True))
    return out


class _Interpreter:
    def __init__(self, scene, base_dir):
        self.scene, self.base_dir = scene, base_dir
        self.vars = {}
        self.funcs = {
            "make_float3": lambda a, b, c: Float3(a, b, c),
            "DiffuseMaterial": lambda c: ScriptMaterial(self._f3(c)),
            "GameObject": lambda m, mat=0xFFFFFFFF: ScriptObject(m, mat),
            "Plane": lambda n, d, m: ScriptPlane(self._f3(n), d, m),
            "scene_add_material": lambda m: int(scene.addMaterial(self._of(m, ScriptMaterial).record())),
            "scene_add_model": self._add_model,
            "scene_add_plane": lambda p: scene.addPlane(self._of(p, ScriptPlane).normal.tuple(), p.d, p.material),
            "scene_add_object": lambda o: int(scene.addObject(self._of(o, ScriptObject).model_id, o.material_id, o.position.tuple(),
                                                                 o.rotation.tuple(), o.scale.tuple())),
        }

    @staticmethod
    def _of(v, cls):
        if not isinstance(v, cls):
            raise ScriptError("expected a %s, got %r" % (cls.__name__, v))
        return v

    def _f3(self, v):
        return self._of(v, Float3)

    def _add_model(self, filename, scale, rotation, offset, material_id, use_mtl):   # Scene::addModel, scene.h:159
        path = filename if os.path.isabs(filename) else os.path.join(self.base_dir, filename)
        return int(self.scene.addModel(path, float(scale), self._f3(rotation).tuple(), self._f3(offset).tuple(), int(material_id), bool(use_mtl)))

    # ---- statements ---------------------------------------------------------------------------------------------
    def run(self, toks):
        self.t, self.i = toks, 0
        while self.t[self.i][0] != "end":
            if self._peek(";"):
                self.i += 1
                continue
            self._statement()

    def _peek(self, op):
        k, v, _ = self.t[self.i]
        return k == "op" and v == op

    def _expect(self, op):
        if not self._peek(op):
            raise ScriptError("expected %r, found %r" % (op, self.t[self.i][1]))
        self.i += 1

    def _statement(self):
        k, v, _ = self.t[self.i]
        if k == "id" and v in ("var", "auto", "global"):
            name = self.t[self.i + 1]
            if name[0] != "id":
                raise ScriptError("expected a name after %r" % v)
            self.i += 2
            self._expect("=")
            self.vars[name[1]] = self._expr()
            return
        j = self.i + 1                                           # assignment  name(.field)* = expr  ?  (decided before anything runs)
        while self.t[j][0] == "op" and self.t[j][1] == "." and self.t[j + 1][0] == "id":
            j += 2
        if k == "id" and self.t[j][0] == "op" and self.t[j][1] == "=":
            target = self._postfix(lvalue=True)
            self.i += 1
            value = self._expr()
            owner, field = target
            if owner is None:
                if field not in self.vars:
                    raise ScriptError("assignment to the undeclared variable %r" % field)
                self.vars[field] = value
            else:
                self._assign_field(owner, field, value)
            return
        self._expr()

    @staticmethod
    def _fields(owner):
        if isinstance(owner, Float3):
            return ("x", "y", "z")
        return getattr(owner, "FIELDS", ())

    def _assign_field(self, owner, field, value):
        if field not in self._fields(owner):
            raise ScriptError("%s has no field %r" % (type(owner).__name__, field))
        cur = getattr(owner, field)
        if isinstance(cur, Float3):
            setattr(owner, field, self._f3(value).copy())      # chai_float_assign: float3 = float3 copies (sceneBuilder.h:259-261,304)
        else:
            setattr(owner, field, float(value))

    # ---- expressions: comparison-free arithmetic with the usual precedence ------------------------------------------
    def _expr(self):
        v = self._term()
        while self._peek("+") or self._peek("-"):
            op = self.t[self.i][1]
            self.i += 1
            r = self._term()
            v = v + r if op == "+" else v - r
        return v

    def _term(self):
        v = self._unary()
        while self._peek("*") or self._peek("/"):
            op = self.t[self.i][1]
            self.i += 1
            r = self._unary()
            if op == "*":
                v = v * r
            elif isinstance(v, int) and isinstance(r, int):
                v = int(v / r)                                               # C++ integer division
            else:
                v = v / r
        return v

    def _unary(self):
        if self._peek("-"):
            self.i += 1
            return -self._unary()
        if self._peek("+"):
            self.i += 1
            return self._unary()
        v = self._postfix(lvalue=False)
        return v

    def _postfix(self, lvalue):
        """primary followed by .field / (args).  With lvalue=True a trailing variable or field is returned as (owner, name)."""
        k, v, _ = self.t[self.i]
        if k == "num":
            self.i += 1
            return int(v) if v.isdigit() else float(v.rstrip("f"))        # integer literals stay integers (1/2 == 0, as in ChaiScript)
        if k == "str":
            self.i += 1
            return bytes(v[1:-1], "utf-8").decode("unicode_escape")
        if k == "op" and v == "(":
            self.i += 1
            r = self._expr()
            self._expect(")")
            return r
        if k != "id":
            raise ScriptError("unexpected %r" % v)
        self.i += 1
        if v in ("true", "false"):
            return v == "true"
        if self._peek("("):
            if v not in self.funcs:
                raise ScriptError("unknown function %r" % v)                 # ChaiScript: "Can not find object"
            cur, ref = self.funcs[v](*self._args()), None
        else:
            ref = (None, v)
            cur = None
        while True:
            if self._peek("."):
                if ref is not None:
                    cur = self._load(ref)
                name = self.t[self.i + 1]
                if name[0] != "id":
                    raise ScriptError("expected a field name after '.'")
                self.i += 2
                ref = (cur, name[1])
                continue
            break
        if ref is None:
            return cur
        if lvalue and self._peek("="):
            return ref
        return self._load(ref)

    def _load(self, ref):
        owner, name = ref
        if owner is None:
            if name not in self.vars:
                raise ScriptError("unknown name %r" % name)
            return self.vars[name]
        if name not in self._fields(owner):
            raise ScriptError("%s has no field %r" % (type(owner).__name__, name))
        return getattr(owner, name)

    def _args(self):
        self._expect("(")
        args = []
        if not self._peek(")"):
            args.append(self._expr())
            while self._peek(","):
                self.i += 1
                args.append(self._expr())
        self._expect(")")
        return args


def run_scene_script(text, base_dir=".", scene=None):
    """Evaluates a scene script against a Scene and finalizes it (getScriptedScene, sceneBuilder.h:263-311)."""
    scene = scene or Scene()
    try:
        _Interpreter(scene, base_dir).run(_tokenize(text))
    except TypeError as e:                                   # wrong argument count / types in a bound call
        raise ScriptError(str(e))
    return scene.finalize()


def load_scene_script(path):
    with open(path) as f:
        return run_scene_script(f.read(), os.path.dirname(os.path.abspath(path)))


# ---- camera state file, stateLoader.h:30-84 ---------------------------------------------------------------------------
INITIAL_CAMERA = ((0.0, 2.0, -3.0), (0.0, 0.0, 1.0), 1.5, 5.0, 0.01)        # initialCamera(), stateLoader.h:30-33


def initial_camera(width, height):
    eye, view, d, focal, aperture = INITIAL_CAMERA
    return make_camera(eye, view, d, focal, aperture, width, height)


def _float3_line(line):                                      # parseFloat3: number, any one character, number, ...
    p = re.split(r"[^0-9eE+\-.]", line.strip())
    p = [x for x in p if x != ""]
    return tuple(float(x) for x in (p + ["0", "0", "0"])[:3])


def read_state(path, width, height):
    """readState(): the saved camera, or initialCamera() when the file cannot be opened."""
    try:
        with open(path) as f:
            lines = f.read().splitlines()
    except OSError:
        return initial_camera(width, height)
    lines += [""] * 5
    num = [np.float32(x.strip() or 0) for x in lines[2:5]]
    return make_camera(_float3_line(lines[0]), _float3_line(lines[1]), float(num[0]), float(num[1]), float(num[2]), width, height)


def save_state(path, camera):
    """saveState(): five lines, floats printed like a default-precision ostream (%g)."""
    c = np.asarray(camera).reshape(-1)[0]
    with open(path, "w") as f:
        f.write("|".join("%g" % x for x in c["eye"]) + "\n")
        f.write("|".join("%g" % x for x in c["viewDir"]) + "\n")
        for k in ("d", "focalLength", "aperture"):
            f.write("%g\n" % c[k])
