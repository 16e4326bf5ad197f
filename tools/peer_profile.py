#!/usr/bin/env python
"""Training frames with the peer-memory exchange kernels on ONE GPU (world 1: the owner range is every triangle and all
traffic is local HBM), for an ncu capture of k_peer_arrive / k_peer_reduce_propagate / k_peer_finish and of k_tail in
pair mode (1/8 of the 1080p image => the tail queue is short).  usage: tools/peer_profile.py [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_pathtracer_b200 import abi, api, scenes
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 6
arrays = scenes.sibenik_scene().arrays()
# rank 0 of an emulated 8-rank job renders its 1/8 of the tiles; the exchange runs over this rank alone
cfg = abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1, rank=0, world_size=8)
pt = api.Pathtracer(arrays, cfg).Init()
cfg1 = abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1)
pt1 = api.Pathtracer(arrays, cfg1).Init()
pt1.peer_connect([pt1.peer_export()])
cams = scenes.camera_sweep("sibenik", 1920, 1080, 16)
for p in (pt, pt1):
    p.Render(cams[0], shouldClear=True)
    for i in range(frames):
        p.Render(cams[i % 16])
    p.Finish()
print("ok", pt.stats().extend_rays, pt1.stats().extend_rays)
pt.close(); pt1.close()
