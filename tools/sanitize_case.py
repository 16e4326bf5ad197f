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
print("ok", float(img[..., 3].min()), np.isfinite(img).all(), post.shape, tiles.shape, pt.stats().extend_rays)
pt.close()
