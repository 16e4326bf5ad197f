"""Pins the CPU restatement (oracle/pt_oracle.cpp) function by function against the reference's own
HYBRID functions compiled as host code (oracle/_ref, built from /root/reference/src by oracle/Makefile).
Bit-exact on seeded random inputs.  Skipped where oracle/_ref is absent; the golden-vector tests in
test_cpu_abi_and_oracle.py cover the same functions end to end there."""
import ctypes as C

import numpy as np
import pytest

from cuda_pathtracer_b200 import abi
from oracle import oracle_py
from oracle.oracle_py import OracleLib, RefLib, RefScene

pytestmark = [pytest.mark.ref, pytest.mark.skipif(not oracle_py.ref_available(160, 120), reason="oracle/_ref not built")]
F = C.c_float


def fa(*a):
    return (C.c_float * len(a))(*[float(x) for x in a])


@pytest.fixture(scope="module")
def libs():
    ref = RefLib(160, 120).L
    ora = OracleLib.lib()
    for L, p in ((ref, "ref_"), (ora, "ora_")):
        getattr(L, p + "rand").restype = C.c_float
    return ref, ora


def bits(x):
    return np.asarray(x, np.float32).view(np.uint32).tolist()


def test_rng_and_seeds(libs):
    ref, ora = libs
    rng = np.random.default_rng(1)
    for s in rng.integers(0, 2**32, 200, dtype=np.uint64):
        assert ref.ref_wang_hash(C.c_uint(int(s))) == ora.ora_wang_hash(C.c_uint(int(s)))
        a, b = C.c_uint(int(s) | 1), C.c_uint(int(s) | 1)
        for _ in range(4):
            assert bits([ref.ref_rand(C.byref(a))]) == bits([ora.ora_rand(C.byref(b))]) and a.value == b.value
    for x, y, r in rng.integers(0, 160, (100, 3)):
        assert ref.ref_get_seed(int(x), int(y) % 120, int(r)) == ora.ora_get_seed(int(x), int(y) % 120, int(r), 160)


def test_vector_helpers_and_samplers(libs):
    ref, ora = libs
    rng = np.random.default_rng(2)
    for _ in range(300):
        v = rng.normal(size=3) * 10 ** rng.uniform(-3, 3)
        o1, o2 = (C.c_float * 3)(), (C.c_float * 3)()
        ref.ref_normalize(fa(*v), o1); ora.ora_normalize(fa(*v), o2)
        assert bits(o1) == bits(o2)
        n = np.array(o1)
        r0, r1 = rng.random(), rng.random()
        ref.ref_sample_cosine(fa(*n), F(r0), F(r1), o1); ora.ora_sample_cosine(fa(*n), F(r0), F(r1), o2)
        assert bits(o1) == bits(o2)
        u1, u2 = (C.c_float * 2)(), (C.c_float * 2)()
        ref.ref_normal_to_uv(fa(*n), u1); ora.ora_normal_to_uv(fa(*n), u2)
        assert bits(u1) == bits(u2)
        rc = list(rng.uniform(0.1, 2.0, 8))
        rc9 = fa(*rc, sum(rc))
        s1, s2 = C.c_uint(12345 + _), C.c_uint(12345 + _)
        b1, b2, p1, p2 = C.c_int(), C.c_int(), C.c_float(), C.c_float()
        ref.ref_sample_cached(fa(*n), C.byref(s1), rc9, o1, C.byref(b1), C.byref(p1))
        ora.ora_sample_cached(fa(*n), C.byref(s2), rc9, o2, C.byref(b2), C.byref(p2))
        assert bits(o1) == bits(o2) and b1.value == b2.value and bits([p1.value]) == bits([p2.value]) and s1.value == s2.value


def test_reflect_refract(libs):
    ref, ora = libs
    rng = np.random.default_rng(3)
    for i in range(300):
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        if np.dot(d, n) > 0:
            n = -n
        pos = rng.normal(size=3) * 5
        ray = fa(*(rng.normal(size=3)), *d)
        o1, o2 = (C.c_float * 7)(), (C.c_float * 7)()
        ior = float(rng.uniform(1.0, 2.0))
        ref.ref_refract(ray, fa(*n), fa(*pos), F(ior), i % 2, o1); ora.ora_refract(ray, fa(*n), fa(*pos), F(ior), i % 2, o2)
        assert bits(o1) == bits(o2)
        q1, q2 = (C.c_float * 6)(), (C.c_float * 6)()
        ref.ref_reflect(ray, fa(*n), fa(*pos), q1); ora.ora_reflect(ray, fa(*n), fa(*pos), q2)
        assert bits(q1) == bits(q2)


def test_primitive_intersections(libs):
    ref, ora = libs
    rng = np.random.default_rng(4)
    for i in range(500):
        o = rng.normal(size=3) * 3
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        if i % 7 == 0:
            d[i % 3] = 0.0                                     # axis-parallel rays: inf / NaN lanes in the slab test
        ray = fa(*o, *d)
        tri = fa(*(rng.normal(size=9) * 2))
        t1, t2 = (C.c_float * 3)(), (C.c_float * 3)()
        h1, h2 = ref.ref_tri_intersect(ray, tri, t1), ora.ora_tri_intersect(ray, tri, t2)
        assert h1 == h2 and (not h1 or bits(t1) == bits(t2))
        a, b = C.c_float(), C.c_float()
        c = rng.normal(size=3) * 2
        h1, h2 = ref.ref_sphere_intersect(ray, fa(*c), F(1.5), C.byref(a)), ora.ora_sphere_intersect(ray, fa(*c), F(1.5), C.byref(b))
        assert h1 == h2 and (not h1 or bits([a.value]) == bits([b.value]))
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        h1, h2 = ref.ref_plane_intersect(ray, fa(*n), F(-3), C.byref(a)), ora.ora_plane_intersect(ray, fa(*n), F(-3), C.byref(b))
        assert h1 == h2 and (not h1 or bits([a.value]) == bits([b.value]))
        lo = rng.normal(size=3) * 2
        box = fa(*lo, *(lo + rng.random(3) * 3))
        with np.errstate(divide="ignore"):
            inv = fa(*(np.float32(1.0) / np.asarray(d, np.float32)))
        if i % 11 == 0:
            ray_o = fa(lo[0], o[1], o[2])                      # origin exactly on a slab plane
        else:
            ray_o = fa(*o)
        h1, h2 = ref.ref_slab(ray_o, inv, box, C.byref(a)), ora.ora_slab(ray_o, inv, box, C.byref(b))
        assert h1 == h2 and (np.isnan(a.value) and np.isnan(b.value) or bits([a.value]) == bits([b.value]))


def test_multi_instance_traversal_incoherent_rays():
    """traverseTopLevel over a scene with rotated / scaled instances, a sphere and a plane, random rays."""
    ref = RefLib(160, 120)
    rs = RefScene(ref)
    m = rs.add_material()
    mod = rs.add_model_obj("/root/reference/teapot.obj", material=m)
    cube = rs.add_model_obj("/root/reference/cube.obj", material=m)
    rs.add_object(mod, pos=(0, 0, 0))
    rs.add_object(mod, pos=(7, 1, 2), rot=(0.3, 1.0, -0.4))
    rs.add_object(cube, pos=(-5, 2, 1), rot=(0.5, 0.2, 0.1), scale=(0.2, 0.2, 0.2))     # example_scene.chai scale: the mixed-space t quirk
    rs.add_object(cube, pos=(2, 6, -3), rot=(0, 0.7, 0))
    rs.add_sphere((-2, -1, -3), 2, 0)
    rs.add_plane((0, -1, 0), -3, 0)
    rs.finalize()
    a = rs.arrays()
    o = OracleLib(abi.default_config(160, 120))
    desc, keep = abi.make_scene_desc(a)
    o.upload_scene(desc, keep)
    rng = np.random.default_rng(5)
    n = 20000
    rays = np.zeros(n, abi.RAY)
    rays["origin"] = rng.normal(size=(n, 3)) * (6, 3, 6) + (1, 2, 0)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["direction"] = d
    rays["length"] = np.where(rng.random(n) < 0.5, 9999999.0, rng.uniform(0.5, 10, n))
    rays["pixeli"] = np.arange(n)
    for any_hit in (False, True):
        assert rs.trace(rays, any_hit, 2).tobytes() == o.trace(rays, any_hit, 2).tobytes()
    o.close()
