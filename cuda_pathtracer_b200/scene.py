"""Python face of the host-side scene code in libptb200.so (include/ptb200_scene.h).

``Scene`` mirrors the reference's ``Scene`` (reference src/scene.h:120-402): addMaterial / addModel /
addObject / addSphere / addPlane / finalize / update, plus ``interactive_depth`` (scene.h:137).
``make_camera`` mirrors the Camera constructor + update() (src/types.h:609-637).
"""
import ctypes as C

import numpy as np

from . import abi
from .api import PtError, load_library

_F3 = C.c_float * 3


def _f3(v):
    return _F3(*[float(x) for x in v])


def _bind(L):
    if getattr(L, "_pts_bound", False):
        return L
    L.pts_scene_new.restype = C.c_void_p
    L.pts_scene_free.argtypes = [C.c_void_p]
    L.pts_last_error.restype = C.c_char_p
    L.pts_last_error.argtypes = [C.c_void_p]
    L.pts_last_build_ms.restype = C.c_double
    L.pts_last_build_ms.argtypes = [C.c_void_p]
    L.pts_set_build_device.argtypes = [C.c_void_p, C.c_int]
    L.pts_set_build_device.restype = None
    L.pts_add_material.argtypes = [C.c_void_p, C.c_void_p]
    L.pts_add_material.restype = C.c_uint32
    L.pts_set_material.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    L.pts_add_model_obj.argtypes = [C.c_void_p, C.c_char_p, C.c_float, _F3, _F3, C.c_uint32, C.c_int]
    L.pts_add_model_soup.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    L.pts_material_texture_name.argtypes = [C.c_void_p, C.c_uint32, C.c_int]
    L.pts_material_texture_name.restype = C.c_char_p
    L.pts_add_object.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, _F3, _F3, _F3]
    L.pts_add_object.restype = C.c_uint32
    L.pts_set_object.argtypes = [C.c_void_p, C.c_uint32, _F3, _F3, _F3]
    L.pts_add_sphere.argtypes = [C.c_void_p, _F3, C.c_float, C.c_uint32]
    L.pts_add_plane.argtypes = [C.c_void_p, _F3, C.c_float, C.c_uint32]
    L.pts_finalize.argtypes = [C.c_void_p]
    L.pts_update.argtypes = [C.c_void_p]
    L.pts_get_desc.argtypes = [C.c_void_p, C.POINTER(abi.SceneDesc)]
    L.pts_camera_make.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    L._pts_bound = True
    return L


def material(diffuse=(1, 1, 1), specular=(0, 0, 0), emission=(0, 0, 0), reflect=0.0, glossy=0.0, transmit=0.0,
             refractive_index=0.0, absorption=(0, 0, 0)):
    """Material::DIFFUSE(color) with the other fields set afterwards, as the presets do (types.h:48-55)."""
    m = np.zeros(1, abi.MATERIAL)
    m["diffuse_color"], m["specular_color"], m["emission"] = diffuse, specular, emission
    m["reflect"], m["glossy"], m["transmit"], m["refractive_index"] = reflect, glossy, transmit, refractive_index
    m["absorption"] = absorption
    return m


def make_camera(eye, view_dir, d, focal_length, aperture, width, height):
    L = _bind(load_library())
    cam = np.zeros(1, abi.CAMERA)
    a = (C.c_float * 9)(*[float(x) for x in (*eye, *view_dir, d, focal_length, aperture)])
    L.pts_camera_make(a, width, height, abi.ptr(cam))
    return cam


class Scene:
    def __init__(self):
        self.L = _bind(load_library())
        self.h = C.c_void_p(self.L.pts_scene_new())
        self.interactive_depth = 1      # scene.h:137
        self.interactive_samples = 1    # scene.h:138
        self._materials = []

    def __del__(self):
        try:
            if self.h:
                self.L.pts_scene_free(self.h)
                self.h = None
        except Exception:
            pass

    def _ck(self, r, what):
        if r < 0:
            raise PtError("%s failed (%d): %s" % (what, r, (self.L.pts_last_error(self.h) or b"").decode()))
        return r

    def addMaterial(self, m):
        m = np.ascontiguousarray(m)
        return int(self.L.pts_add_material(self.h, abi.ptr(m)))

    def setMaterial(self, idx, m):
        m = np.ascontiguousarray(m)
        self._ck(self.L.pts_set_material(self.h, idx, abi.ptr(m)), "pts_set_material")

    def addModel(self, filename, scale=1.0, rotation=(0, 0, 0), offset=(0, 0, 0), material=0, useMtl=False):
        return self._ck(self.L.pts_add_model_obj(self.h, filename.encode(), scale, _f3(rotation), _f3(offset), material, int(useMtl)),
                        "pts_add_model_obj")

    def addModelSoup(self, verts, normals=None, uvs=None, mats=None, material=0):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        normals = None if normals is None else np.ascontiguousarray(normals, np.float32)
        uvs = None if uvs is None else np.ascontiguousarray(uvs, np.float32)
        mats = None if mats is None else np.ascontiguousarray(mats, np.uint32)
        return self._ck(self.L.pts_add_model_soup(self.h, abi.ptr(verts), abi.ptr(normals), abi.ptr(uvs), abi.ptr(mats), len(verts), material),
                        "pts_add_model_soup")

    def textureName(self, material_id, normal_map=False):
        return (self.L.pts_material_texture_name(self.h, material_id, int(normal_map)) or b"").decode()

    def addObject(self, model, material=0xFFFFFFFF, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1)):
        return int(self.L.pts_add_object(self.h, model, material, _f3(position), _f3(rotation), _f3(scale)))

    def setObject(self, idx, position, rotation, scale):
        self._ck(self.L.pts_set_object(self.h, idx, _f3(position), _f3(rotation), _f3(scale)), "pts_set_object")

    def addSphere(self, pos, radius, material):
        self.L.pts_add_sphere(self.h, _f3(pos), radius, material)

    def addPlane(self, normal, d, material):
        self.L.pts_add_plane(self.h, _f3(normal), d, material)

    def finalize(self):
        self._ck(self.L.pts_finalize(self.h), "pts_finalize")
        return self

    def update(self):
        self._ck(self.L.pts_update(self.h), "pts_update")

    def set_build_device(self, device):
        """device >= 0: models added from now on get their BVH2 from pt_build_bvh on that GPU; -1: host builder."""
        self.L.pts_set_build_device(self.h, int(device))
        return self

    def last_build_ms(self):
        return float(self.L.pts_last_build_ms(self.h))

    def desc(self):
        d = abi.SceneDesc()
        self._ck(self.L.pts_get_desc(self.h, C.byref(d)), "pts_get_desc")
        return d

    def arrays(self):
        """numpy copies of the host vectors (same keys as abi.make_scene_desc)."""
        d = self.desc()

        def arr(p, n, dt):
            if not p or n == 0:
                return np.zeros(0, dt)
            return np.frombuffer((C.c_char * (n * dt.itemsize)).from_address(p), dtype=dt, count=n).copy()
        a = {"vertices": arr(d.vertices, d.num_triangles, abi.TRIANGLE_V), "vertexData": arr(d.vertexData, d.num_triangles, abi.TRIANGLE_D),
             "models": arr(d.models, d.num_models, abi.MODEL), "materials": arr(d.materials, d.num_materials, abi.MATERIAL),
             "instances": arr(d.instances, d.num_instances, abi.INSTANCE), "topBvh": arr(d.topBvh, d.num_tlas_nodes, abi.TLAS_NODE),
             "spheres": arr(d.spheres, d.num_spheres, abi.SPHERE), "planes": arr(d.planes, d.num_planes, abi.PLANE)}
        a["bvh"] = [arr(int(m["bvh"]), int(m["nrBvhNodes"]), abi.BVH_NODE) for m in a["models"]]
        return a
