#!/usr/bin/env python
"""Small end-to-end case for compute-sanitizer: every kernel of the path at 96x64 (wavefront, pair, tail, guiding, post)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from cuda_pathtracer_b200 import abi, api, scenes
a = scenes.sibenik_scene().arrays()
cfg = abi.default_config(96, 64, max_ray_depth=8, interactive_depth=1)
pt = api.Pathtracer(a, cfg).Init()
pt.set_blue_noise(np.random.default_rng(1).random((16, 16)).astype(np.float32))
cam = scenes.camera_sweep("sibenik", 96, 64, 16)[0]
pt.Render(cam, shouldClear=True)
for _ in range(6): pt.Render(cam)
pt.Finish()
img = pt.read_image(); post = pt.postprocess(True); tiles = pt.read_tiles()
# round 2: cooperative traversal + mixed pair work list (a lone training frame on a small image), ray sorting, sky sampling,
# the static-scratch trace entry, the host framebuffer ring (one rank), the counting kernels
from cuda_pathtracer_b200.distributed import HostRing
pt.set_skydome(np.ones((16, 32, 4), np.float32)); pt.set_ray_sorting(True, 3, 2); pt.set_sky_sampling(True)
pt.HCACHE = True
pt.Render(cam, shouldClear=True)
for _ in range(4): pt.Render(cam)
pt.set_ray_sorting(False)
rays = pt.debug_primary_rays(cam, 0)
h0, h1 = pt.debug_trace(rays, False), pt.trace_static(rays, False)
assert np.array_equal(h0["primitive_id"], h1["primitive_id"])
ring = HostRing(pt, slots=2)
for _ in range(3):
    pt.Render(cam); e = pt.deliver_image_async(); v = pt.host_ring_wait(e); assert np.isfinite(v).all(); pt.host_ring_release(e)
ring.close()
pt.set_counting(True); pt.Render(cam); pt.Finish(); print("tally", pt.tally()); pt.set_counting(False)
print("ok", float(img[..., 3].min()), np.isfinite(img).all(), post.shape, tiles.shape, pt.stats().extend_rays)
pt.close()
