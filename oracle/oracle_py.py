"""TEST INFRASTRUCTURE -- ctypes bindings for the two CPU checkers.

* ``OracleLib``  -> oracle/libptoracle.so  (oracle/pt_oracle.cpp, the CPU restatement; always buildable)
* ``RefLib``     -> oracle/_ref/libptref_WxH.so (the reference's own headers compiled as host code,
                    oracle/ref_harness.cu; only present when built in a container that has /root/reference)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from cuda_pathtracer_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))
F = C.c_float


def _fa(*a):
    return (C.c_float * len(a))(*[float(x) for x in a])


def build_oracle(force=False):
    so = os.path.join(HERE, "libptoracle.so")
    src = os.path.join(HERE, "pt_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "libptoracle.so"], stdout=subprocess.DEVNULL)
    return so


def ref_available(w=640, h=480):
    return os.path.exists(os.path.join(HERE, "_ref", "libptref_%dx%d.so" % (w, h)))


class OracleLib:
    """One CPU-restatement renderer (mirrors the pt_* C ABI one-to-one)."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = C.CDLL(build_oracle())
            L.ora_create.restype = C.c_void_p
            L.ora_add_texture.restype = C.c_uint64
            L.ora_rand.restype = C.c_float
            for n in ("ora_destroy", "ora_set_threads", "ora_upload_scene", "ora_update_instances", "ora_set_toggles",
                      "ora_set_skydome", "ora_set_sky_sampling", "ora_set_sky_constant", "ora_set_blue_noise", "ora_add_texture", "ora_render",
                      "ora_read_image", "ora_read_radiance_cache", "ora_get_stats", "ora_reset_stats",
                      "ora_get_rand_state", "ora_primary_rays", "ora_trace", "ora_trace_stack_depth"):
                getattr(L, n).argtypes = None
            cls._lib = L
        return cls._lib

    def __init__(self, cfg):
        self.L = self.lib()
        self.cfg = cfg
        self.h = C.c_void_p(self.L.ora_create(C.byref(cfg)))
        self._keep = []

    def close(self):
        if self.h:
            self.L.ora_destroy(self.h)
            self.h = None

    def set_threads(self, n):
        self.L.ora_set_threads(self.h, C.c_int(n))

    def upload_scene(self, desc, keep=None):
        self._keep.append((desc, keep))
        return self.L.ora_upload_scene(self.h, C.byref(desc))

    def update_instances(self, instances, tlas):
        self._keep.append((instances, tlas))
        self.L.ora_update_instances(self.h, abi.ptr(instances), C.c_uint32(len(instances)), abi.ptr(tlas), C.c_uint32(len(tlas)))

    def set_toggles(self, nee, guiding):
        self.L.ora_set_toggles(self.h, C.c_int(int(nee)), C.c_int(int(guiding)))

    def set_skydome(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        self._keep.append(rgba)
        self.L.ora_set_skydome(self.h, abi.ptr(rgba), C.c_uint32(rgba.shape[1]), C.c_uint32(rgba.shape[0]))

    def set_sky_constant(self, r, g, b):
        self.L.ora_set_sky_constant(self.h, F(r), F(g), F(b))

    def set_sky_sampling(self, on):
        if self.L.ora_set_sky_sampling(self.h, C.c_int(int(on))) != 0:
            raise RuntimeError("ora_set_sky_sampling: no skydome set")

    def set_blue_noise(self, lum):
        lum = np.ascontiguousarray(lum, np.float32)
        self._keep.append(lum)
        self.L.ora_set_blue_noise(self.h, abi.ptr(lum), C.c_uint32(lum.shape[1]), C.c_uint32(lum.shape[0]))

    def add_texture(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        self._keep.append(rgba)
        return int(self.L.ora_add_texture(self.h, abi.ptr(rgba), C.c_uint32(rgba.shape[1]), C.c_uint32(rgba.shape[0])))

    def postprocess(self, lum, albedo, nr_samples, blur=True):
        """CPU restatement of the reference's display shaders (main.cpp:46-171): float4 images in, RGBA out."""
        lum = np.ascontiguousarray(lum, np.float32); albedo = np.ascontiguousarray(albedo, np.float32)
        out = np.empty_like(lum)
        self.L.ora_postprocess(abi.ptr(lum), abi.ptr(albedo), C.c_int(lum.shape[1]), C.c_int(lum.shape[0]), C.c_float(nr_samples), C.c_int(int(blur)), abi.ptr(out))
        return out

    def render(self, cam, should_clear):
        self.L.ora_render(self.h, abi.ptr(cam), C.c_int(int(should_clear)))

    def read_image(self, which=0):
        out = np.zeros((self.cfg.height, self.cfg.width, 4), np.float32)
        self.L.ora_read_image(self.h, abi.ptr(out), C.c_int(which))
        return out

    def read_radiance_cache(self, n):
        out = np.zeros(n, abi.RADIANCE_CACHE)
        self.L.ora_read_radiance_cache(self.h, abi.ptr(out))
        return out

    def stats(self):
        st = abi.Stats()
        cnt = (C.c_uint64 * 6)()
        self.L.ora_get_stats(self.h, C.byref(st), cnt)
        return st, list(cnt)

    def reset_stats(self):
        self.L.ora_reset_stats(self.h)

    def primary_rays(self, cam, rand_idx):
        out = np.zeros(self.cfg.width * self.cfg.height, abi.RAY)
        self.L.ora_primary_rays(self.h, abi.ptr(cam), C.c_uint32(rand_idx), abi.ptr(out))
        return out

    def trace(self, rays, any_hit=False, threads=0, counters=None):
        out = np.zeros(len(rays), abi.HIT)
        cnt = (C.c_uint64 * 3)()
        self.L.ora_trace(self.h, abi.ptr(rays), C.c_uint32(len(rays)), C.c_int(int(any_hit)), abi.ptr(out), C.c_int(threads), cnt)
        if counters is not None:
            counters[:] = list(cnt)
        return out


def oracle_stack_depth(ora, rays):
    """Deepest bottom-level BVH2 stack each ray needs in the reference's traversal order (its array holds 18 entries, kernels.h:207)."""
    rays = np.ascontiguousarray(rays)
    out = np.zeros(len(rays), np.uint32)
    ora.L.ora_trace_stack_depth(ora.h, abi.ptr(rays), C.c_uint32(len(rays)), abi.ptr(out))
    return out


def oracle_camera(eye, view_dir, d, focal, aperture, w, h):
    """Camera as the reference host prepares it (ctor + update(), types.h:609-637), restated."""
    L = OracleLib.lib()
    cam = np.zeros(1, abi.CAMERA)
    L.ora_camera_make(_fa(*eye, *view_dir, d, focal, aperture), C.c_uint32(w), C.c_uint32(h), abi.ptr(cam))
    return cam


class RefLib:
    """The reference's own host code (oracle/_ref).  One library per compile-time resolution."""

    def __init__(self, w=640, h=480):
        path = os.path.join(HERE, "_ref", "libptref_%dx%d.so" % (w, h))
        self.L = L = C.CDLL(path, mode=os.RTLD_LOCAL)
        self.w, self.h = w, h
        L.ref_scene_new.restype = C.c_void_p
        L.ref_ptr.restype = C.c_void_p
        L.ref_rand.restype = C.c_float
        L.ref_time_bvh_build.restype = C.c_double
        assert L.ref_width() == w and L.ref_height() == h

    def sizes(self):
        out = (C.c_uint * 18)()
        self.L.ref_sizes(out)
        return list(out)

    def camera(self, eye, view_dir, d, focal, aperture):
        out = (C.c_float * 18)()
        self.L.ref_camera(_fa(*eye, *view_dir, d, focal, aperture), out)
        o = np.array(out, np.float32)
        cam = np.zeros(1, abi.CAMERA)
        for i, k in enumerate(("eye", "viewDir", "lt", "u", "v")):
            cam[0][k] = o[i * 3:i * 3 + 3]
        cam[0]["d"], cam[0]["focalLength"], cam[0]["aperture"] = o[15], o[16], o[17]
        return cam

    def primary_rays(self, eye, view_dir, d, focal, aperture, rand_idx):
        out = np.zeros(self.w * self.h, abi.RAY)
        self.L.ref_primary_rays(_fa(*eye, *view_dir, d, focal, aperture), C.c_uint(rand_idx), abi.ptr(out))
        return out


class RefScene:
    """Scene built by the reference's own Scene / createBVHBinned / BuildTopLevelBVH (scene.h, bvhBuilder.h)."""

    def __init__(self, ref):
        self.ref = ref
        self.L = ref.L
        self.h = C.c_void_p(self.L.ref_scene_new())

    def add_material(self, diffuse=(1, 1, 1), specular=(0, 0, 0), emission=(0, 0, 0), reflect=0, glossy=0, transmit=0, ior=0,
                     absorption=(0, 0, 0)):
        return self.L.ref_add_material(self.h, _fa(*diffuse, *specular, *emission, reflect, glossy, transmit, ior, *absorption))

    def add_model_obj(self, path, scale=1.0, rot=(0, 0, 0), off=(0, 0, 0), material=0, use_mtl=False):
        return self.L.ref_add_model_obj(self.h, path.encode(), F(scale), _fa(*rot), _fa(*off), C.c_uint(material), C.c_int(int(use_mtl)))

    def add_model_soup(self, verts, normals=None, uvs=None, mats=None, material=0):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = len(verts)
        normals = None if normals is None else np.ascontiguousarray(normals, np.float32)
        uvs = None if uvs is None else np.ascontiguousarray(uvs, np.float32)
        mats = None if mats is None else np.ascontiguousarray(mats, np.uint32)
        return self.L.ref_add_model_soup(self.h, abi.ptr(verts), abi.ptr(normals), abi.ptr(uvs), abi.ptr(mats), C.c_uint(n), C.c_uint(material))

    def add_object(self, model, material=0xFFFFFFFF, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1)):
        self.L.ref_add_object(self.h, C.c_uint(model), C.c_uint(material), _fa(*pos), _fa(*rot), _fa(*scale))

    def set_object(self, idx, pos, rot, scale):
        self.L.ref_set_object(self.h, C.c_uint(idx), _fa(*pos), _fa(*rot), _fa(*scale))

    def add_sphere(self, pos, radius, material):
        self.L.ref_add_sphere(self.h, _fa(*pos), F(radius), C.c_uint(material))

    def add_plane(self, normal, d, material):
        self.L.ref_add_plane(self.h, _fa(*normal), F(d), C.c_uint(material))

    def finalize(self):
        self.L.ref_finalize(self.h)
        self.L.ref_update(self.h)

    def update(self):
        self.L.ref_update(self.h)

    def counts(self):
        out = (C.c_uint * 7)()
        self.L.ref_counts(self.h, out)
        return dict(zip(("triangles", "models", "instances", "tlas", "materials", "spheres", "planes"), out))

    def _arr(self, what, dtype, n, idx=0):
        p = self.L.ref_ptr(self.h, C.c_int(what), C.c_uint(idx))
        if n == 0 or not p:
            return np.zeros(0, dtype)
        buf = (C.c_char * (n * dtype.itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dtype, count=n).copy()

    def arrays(self):
        """Copies of every array the reference's Pathtracer::Init would upload (pathtracer.h:129-190)."""
        c = self.counts()
        a = {
            "vertices": self._arr(0, abi.TRIANGLE_V, c["triangles"]),
            "vertexData": self._arr(1, abi.TRIANGLE_D, c["triangles"]),
            "models": self._arr(2, abi.MODEL, c["models"]),
            "instances": self._arr(3, abi.INSTANCE, c["instances"]),
            "topBvh": self._arr(4, abi.TLAS_NODE, c["tlas"]),
            "materials": self._arr(5, abi.MATERIAL, c["materials"]),
            "spheres": self._arr(6, abi.SPHERE, c["spheres"]),
            "planes": self._arr(7, abi.PLANE, c["planes"]),
        }
        a["bvh"] = [self._arr(8, abi.BVH_NODE, int(m["nrBvhNodes"]), i) for i, m in enumerate(a["models"])]
        return a

    def trace(self, rays, any_hit=False, threads=0):
        out = np.zeros(len(rays), abi.HIT)
        self.L.ref_trace(self.h, abi.ptr(rays), C.c_uint(len(rays)), C.c_int(int(any_hit)), abi.ptr(out), C.c_int(threads))
        return out
