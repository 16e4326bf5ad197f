"""ctypes / numpy mirrors of the POD structs in include/ptb200.h.

Every struct is layout-compatible with the reference struct named beside it (reference paths are
relative to /root/reference/); tests/test_abi.py pins sizes and offsets against the C header and,
when oracle/_ref is built, against the reference's own headers.
"""
import ctypes as C

import numpy as np

# ---- numpy dtypes (arrays of these are handed to the C ABI as-is) -----------------------------------
F3 = [("x", "<f4"), ("y", "<f4"), ("z", "<f4")]


def _dt(fields, itemsize):
    names, formats, offsets = zip(*fields)
    return np.dtype({"names": list(names), "formats": list(formats), "offsets": list(offsets), "itemsize": itemsize})


# Material, src/types.h:33-56
MATERIAL = _dt([
    ("diffuse_color", ("<f4", 3), 0), ("specular_color", ("<f4", 3), 12), ("emission", ("<f4", 3), 24),
    ("reflect", "<f4", 36), ("glossy", "<f4", 40), ("transmit", "<f4", 44), ("refractive_index", "<f4", 48),
    ("absorption", ("<f4", 3), 52), ("texture", "<u8", 64), ("normal_texture", "<u8", 72),
    ("hasTexture", "u1", 80), ("hasNormalMap", "u1", 81)], 88)
SPHERE = _dt([("pos", ("<f4", 3), 0), ("radius", "<f4", 12), ("material", "<u4", 16)], 20)          # types.h:58-69
PLANE = _dt([("normal", ("<f4", 3), 0), ("d", "<f4", 12), ("material", "<u4", 16)], 20)             # types.h:124-133
TRIANGLE_V = _dt([("v0", ("<f4", 3), 0), ("v1", ("<f4", 3), 12), ("v2", ("<f4", 3), 24)], 48)       # types.h:79-83
TRIANGLE_D = _dt([("normal", ("<f4", 3), 0), ("tangent", ("<f4", 3), 12), ("bitangent", ("<f4", 3), 24),
                  ("uv0", ("<f4", 2), 40), ("uv1", ("<f4", 2), 48), ("uv2", ("<f4", 2), 56), ("material", "<u4", 64)], 80)  # types.h:85-96
RADIANCE_CACHE = _dt([("radianceCache", ("<f4", 8), 0), ("additionCache", ("<f4", 8), 32),
                      ("additionCacheCount", ("<f4", 8), 64), ("radianceTotal", "<f4", 96)], 112)   # types.h:98-104
TRIANGLE_LIGHT = _dt([("triangle_index", "<u4", 0), ("instance_index", "<u4", 4)], 8)               # types.h:106-110
RAY = _dt([("origin", ("<f4", 3), 0), ("length", "<f4", 12), ("direction", ("<f4", 3), 16), ("pixeli", "<u4", 28)], 32)  # RayPacked, types.h:273-294
HIT = _dt([("primitive_type", "<u4", 0), ("primitive_id", "<u4", 4), ("instance_id", "<u4", 8), ("t", "<f4", 12)], 16)   # HitInfo, types.h:299-308
SAMPLE_CACHE = _dt([("sample_type", "<u4", 0), ("triangle_id", "<u4", 4), ("cache_bucket_id", "<i4", 8),
                    ("cum_mask", ("<f4", 3), 12)], 32)                                              # types.h:339-345
BVH_NODE = _dt([("vmin", ("<f4", 3), 0), ("first", "<u4", 12), ("vmax", ("<f4", 3), 16), ("count", "<u4", 28)], 32)      # types.h:347-375
MODEL = _dt([("bvh", "<u8", 0), ("triangleStart", "<u4", 8), ("nrTriangles", "<u4", 12), ("nrBvhNodes", "<u4", 16)], 32)  # types.h:377-383
INSTANCE = _dt([("model_id", "<u4", 0), ("transform", ("<f4", (3, 4)), 16), ("invTransform", ("<f4", (3, 4)), 64),
                ("material_id", "<u4", 112)], 128)                                                  # types.h:408-414
TLAS_NODE = _dt([("vmin", ("<f4", 3), 0), ("vmax", ("<f4", 3), 12), ("children", "<u4", 24), ("leaf", "<u4", 28)], 32)   # types.h:431-459
CAMERA = _dt([("_has_moved", "<u4", 0), ("eye", ("<f4", 3), 4), ("viewDir", ("<f4", 3), 16), ("lt", ("<f4", 3), 28),
              ("u", ("<f4", 3), 40), ("v", ("<f4", 3), 52), ("d", "<f4", 64), ("focalLength", "<f4", 68),
              ("aperture", "<f4", 72)], 76)                                                         # types.h:586-677


# ---- ctypes structs passed by pointer ------------------------------------------------------------------
class SceneDesc(C.Structure):          # pt_scene_desc
    _fields_ = [("vertices", C.c_void_p), ("vertexData", C.c_void_p), ("num_triangles", C.c_uint32),
                ("models", C.c_void_p), ("num_models", C.c_uint32),
                ("materials", C.c_void_p), ("num_materials", C.c_uint32),
                ("instances", C.c_void_p), ("num_instances", C.c_uint32),
                ("topBvh", C.c_void_p), ("num_tlas_nodes", C.c_uint32),
                ("spheres", C.c_void_p), ("num_spheres", C.c_uint32),
                ("planes", C.c_void_p), ("num_planes", C.c_uint32)]


class Config(C.Structure):             # pt_config
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("max_ray_depth", C.c_uint32),
                ("interactive_depth", C.c_uint32), ("guiding_frames", C.c_uint32), ("blue_noise_frames", C.c_uint32),
                ("device", C.c_int32), ("rank", C.c_uint32), ("world_size", C.c_uint32), ("flags", C.c_uint32)]


class Stats(C.Structure):              # pt_stats
    _fields_ = [("primary_rays", C.c_uint64), ("extend_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("shade_calls", C.c_uint64), ("samples", C.c_uint64),
                ("wave_extend_rays", C.c_uint64), ("wave_shadow_rays", C.c_uint64), ("kernel_launches", C.c_uint64)]


ALLREDUCE_HOOK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)


FLAG_WAVEFRONT_ONLY, FLAG_NO_PIPELINE, FLAG_HOST_WIDE_BVH, FLAG_NO_PDL, FLAG_SORT_RAYS = 1, 2, 4, 8, 16      # PT_FLAG_* (include/ptb200.h)
HOST_RING_HEADER_BYTES = 4096          # PT_HOST_RING_HEADER_BYTES (include/ptb200.h)
STATIC_TRACE_MAX_RAYS = 307200            # PT_STATIC_TRACE_MAX_RAYS (include/ptb200.h): the static scratch of pt_trace_static


def default_config(width=640, height=480, max_ray_depth=32, interactive_depth=1, guiding_frames=100,
                   blue_noise_frames=1, device=0, rank=0, world_size=1, flags=0):
    """Reference defaults: constants.h:4-14, scene.h:137, pathtracer.h:292, kernels.h:23."""
    return Config(width, height, max_ray_depth, interactive_depth, guiding_frames, blue_noise_frames,
                  device, rank, world_size, flags)


def ptr(a):
    """void* of a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def make_scene_desc(a):
    """pt_scene_desc over a dict of numpy arrays (keys: vertices, vertexData, models, bvh[list], materials,
    instances, topBvh, spheres, planes).  Model.bvh host pointers are patched to the per-model BVH arrays.
    Returns (desc, keepalive) -- keep `keepalive` referenced for as long as the desc is in use."""
    models = a["models"].copy()
    bvh = [np.ascontiguousarray(b) for b in a["bvh"]]
    for i, b in enumerate(bvh):
        models[i]["bvh"] = b.ctypes.data
    keep = {k: np.ascontiguousarray(a[k]) for k in ("vertices", "vertexData", "materials", "instances", "topBvh", "spheres", "planes")}
    keep["models"], keep["bvh"] = models, bvh
    d = SceneDesc()
    d.vertices, d.vertexData, d.num_triangles = ptr(keep["vertices"]), ptr(keep["vertexData"]), len(keep["vertices"])
    d.models, d.num_models = ptr(models), len(models)
    d.materials, d.num_materials = ptr(keep["materials"]), len(keep["materials"])
    d.instances, d.num_instances = ptr(keep["instances"]), len(keep["instances"])
    d.topBvh, d.num_tlas_nodes = ptr(keep["topBvh"]), len(keep["topBvh"])
    d.spheres, d.num_spheres = ptr(keep["spheres"]), len(keep["spheres"])
    d.planes, d.num_planes = ptr(keep["planes"]), len(keep["planes"])
    return d, keep
