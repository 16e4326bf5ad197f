#!/usr/bin/env python
"""A few cache-training frames of rank 0's share of an emulated WORLD-rank job (sibenik 1080p) -- run under
`ncu --metrics gpu__time_duration.sum` to get the device time of every launch of the frame's dependent chain.
usage: tools/shard_launches.py WORLD [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_pathtracer_b200 import abi, api, scenes
world = int(sys.argv[1]); frames = int(sys.argv[2]) if len(sys.argv) > 2 else 3
arrays = scenes.sibenik_scene().arrays()
cfg = abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1, rank=0, world_size=world)
pt = api.Pathtracer(arrays, cfg).Init()
cams = scenes.camera_sweep("sibenik", 1920, 1080, 16)
pt.Render(cams[0], shouldClear=True)
for i in range(frames): pt.Render(cams[i % 16])
pt.Finish()
pt.set_counting(True)
pt.Render(cams[3]); pt.Finish()
print("tally", pt.tally(), "queues", [list(map(int, x)) for x in pt.debug_read_counters(8)])
pt.close()
