#!/usr/bin/env python
"""How long does ONE short traversal launch take, and what bounds it?  Incoherent rays inside the sibenik proxy (origins on
bounce-1 hit points, random directions), queue sizes from 64 to 256k, one-lane and cooperative form; next to the time, the
launch's work as the counting kernel sees it.  usage: tools/latency_probe.py"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from cuda_pathtracer_b200 import abi, api, scenes

arrays = scenes.sibenik_scene().arrays()
cam = scenes.camera_sweep("sibenik", 1920, 1080, 16)[3]


def make(coop):
    os.environ["PTB200_COOP"] = coop
    cfg = abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1)
    return api.Pathtracer(arrays, cfg).Init()


pt = make("4,2")
prim = pt.debug_primary_rays(cam, 0) if hasattr(pt, "debug_primary_rays") else None
hits = pt.debug_trace(prim, False)
ok = hits["primitive_id"] != 0xFFFFFFFF
rng = np.random.default_rng(1)
o = prim["origin"][ok] + prim["direction"][ok] * (hits["t"][ok, None] * np.float32(0.999))
d = rng.normal(size=o.shape).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
pool = np.zeros(len(o), abi.RAY)
pool["origin"], pool["length"], pool["direction"], pool["pixeli"] = o, np.float32(9999999.0), d, np.arange(len(o), dtype=np.uint32)
pool = pool[rng.permutation(len(pool))]
pt.close()
for coop in ("0", "4,2", "100000"):
    pt = make(coop)
    for n in (64, 1024, 4096, 20000, 63000, 250000, 1000000):
        rays = pool[:n]
        for any_hit in (False, True):
            _, ms = pt.debug_trace_timed(rays, any_hit, repeats=6)
            print(json.dumps({"coop": coop, "n": n, "any": any_hit, "us": round(ms * 1e3, 1), "mrays_s": round(n / ms / 1e3, 1)}), flush=True)
    pt.close()
