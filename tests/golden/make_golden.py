"""Generates the committed golden fixtures from oracle/_ref (the reference's own host code compiled
from /root/reference/src).  Run in the build container only:  python tests/golden/make_golden.py

  teapot_scene.npz   every array Pathtracer::Init would upload for config C1 (teapot.obj, one identity
                     instance, Material::DIFFUSE(1)) after the reference's addModel/createBVHBinned/
                     BuildTopLevelBVH, plus the reference's primary rays + closest/any hits at 160x120
  outside_scene.npz  getOutsideScene (sceneBuilder.h:15-117): 10 glass cube instances + checker plane
  known_answers.json the 640x480 teapot known-answer numbers of BASELINE.md section 4
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle_py import RefLib, RefScene  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CAM_TEAPOT = ((0, 2, -6), (0, 0, 1), 1.5, 5, 0.01)
CAM_OUTSIDE = ((9.81771, 13.5924, -1.12687), (-0.65847, -0.544656, 0.519391), 1.5, 8.49543, 6.50994e-10)   # save.txt:1-5


def obj_order_soup(path):
    """Triangle vertices in OBJ file order (before any BVH permutation), float32 (n, 9)."""
    v, tris = [], []
    for line in open(path):
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            v.append([np.float32(float(x)) for x in t[1:4]])
        elif t[0] == "f":
            idx = [int(x.split("/")[0]) - 1 for x in t[1:]]
            assert len(idx) == 3
            tris.append(v[idx[0]] + v[idx[1]] + v[idx[2]])
    return np.array(tris, np.float32) + np.float32(0.0)   # the identity bake-in transform (scene.h:285-291) turns -0 into +0


def pack(arrays):
    out = {k: v for k, v in arrays.items() if k != "bvh"}
    for i, b in enumerate(arrays["bvh"]):
        out["bvh%d" % i] = b
    return out


def teapot(ref):
    s = RefScene(ref)
    m = s.add_material()
    mod = s.add_model_obj("/root/reference/teapot.obj", material=m)
    s.add_object(mod)
    s.finalize()
    return s


def outside(ref):
    s = RefScene(ref)
    s.add_material(diffuse=(0.4, 0.4, 0.4))                                     # white
    cube = s.add_material(diffuse=(1, 1, 1), transmit=1.0, ior=1.1, glossy=0.02, absorption=(0.1, 0.5, 0.8))
    mod = s.add_model_obj("/root/reference/cube.obj", material=cube)
    f = np.float32(0)
    for i in range(10):                                                           # handler at t = 0 (sceneBuilder.h:89-100)
        pos = (float(np.float32(10 * np.sin(np.float32(f + np.float32(0))))), 0.0, float(np.float32(10 * np.cos(np.float32(f)))))
        s.add_object(mod, pos=pos, rot=(float(f), 0, 0))
        f = np.float32(f + np.float32(2 * 0.3141592))
    s.add_plane((0, -1, 0), -3, 0)
    s.finalize()
    return s


def main():
    ref_small = RefLib(160, 120)
    ref_full = RefLib(640, 480)
    for name, build, cam in (("teapot", teapot, CAM_TEAPOT), ("outside", outside, CAM_OUTSIDE)):
        s = build(ref_small)
        a = pack(s.arrays())
        rays = ref_small.primary_rays(*cam, 0)
        a["cam_in"] = np.array([*cam[0], *cam[1], cam[2], cam[3], cam[4]], np.float32)
        a["camera_160x120"] = ref_small.camera(*cam)
        a["rays_160x120"] = rays
        a["hits_160x120"] = s.trace(rays, False, 1)
        a["anyhits_160x120"] = s.trace(rays, True, 1)
        if name == "teapot":
            a["soup_obj_order"] = obj_order_soup("/root/reference/teapot.obj")
        np.savez_compressed(os.path.join(HERE, name + "_scene.npz"), **a)
    s = teapot(ref_full)
    rays = ref_full.primary_rays(*CAM_TEAPOT, 0)
    hits = s.trace(rays, False, 1)
    hit = hits["primitive_id"] != 0xFFFFFFFF
    ka = {"config": "C1 teapot 640x480", "triangles": int(s.counts()["triangles"]), "bvh_nodes": int(s.arrays()["models"][0]["nrBvhNodes"]),
          "hit_pixels": int(hit.sum()), "sum_primitive_id": int(hits["primitive_id"][hit].astype(np.uint64).sum()),
          "sum_t": float(hits["t"][hit].astype(np.float64).sum())}
    json.dump(ka, open(os.path.join(HERE, "known_answers.json"), "w"), indent=1)
    print(ka)


if __name__ == "__main__":
    main()
