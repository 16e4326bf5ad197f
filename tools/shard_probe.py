#!/usr/bin/env python
"""Times one rank's share of the sibenik 1080p workload on ONE GPU (rank 0 of an emulated world size): what a
rank of an N-GPU job does per sample, without the collective.  usage: tools/shard_probe.py WORLD [steps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cuda_pathtracer_b200 import abi, api, scenes
world = int(sys.argv[1]); steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
arrays = scenes.sibenik_scene().arrays()
cfg = abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1, rank=0, world_size=world)
pt = api.Pathtracer(arrays, cfg).Init()
cams = scenes.camera_sweep("sibenik", 1920, 1080, 16)
pt.Render(cams[0], shouldClear=True); pt.Finish()
stream = torch.cuda.ExternalStream(pt.stream())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(5): pt.Render(cams[i % 16])        # training frames (radiance cache update + propagate, no collective here)
pt.join(); e0.record(stream)
for i in range(80): pt.Render(cams[i % 16])
pt.join(); e1.record(stream); pt.Finish()
print("world %d: training %.3f ms/frame" % (world, e0.elapsed_time(e1) / 80))
for i in range(30): pt.Render(cams[i % 16])       # past the guiding-training frames
pt.Finish(); pt.reset_stats()
e0.record(stream); t0 = time.perf_counter()
for i in range(steps): pt.Render(cams[i % 16])
host_us = (time.perf_counter() - t0) / steps * 1e6
pt.join(); e1.record(stream); pt.Finish()
ms = e0.elapsed_time(e1); st = pt.stats()
print("  host time per Render call (enqueue only): %.1f us" % host_us)
print("world %d: %.3f ms/sample, %.1f Mrays/s on this rank (x%d = %.1f)" % (world, ms / steps, (st.extend_rays + st.shadow_rays) / ms / 1e3, world, world * (st.extend_rays + st.shadow_rays) / ms / 1e3))
pt.close()
# per-stage device time of one sample at this shard size (serialised, event pairs around every launch)
pt = api.Pathtracer(arrays, cfg).Init()
pt.Render(cams[0], shouldClear=True)
for i in range(110): pt.Render(cams[i % 16])
pt.Finish(); pt.set_profiling(True)
for i in range(50): pt.Render(cams[i % 16])
pt.Finish()
print("  stage us/sample:", {k: round(1e3 * v[0] / 50, 1) for k, v in pt.profile().items()}, "launches/sample", sum(v[1] for v in pt.profile().values()) / 50)
pt.close()
pt = api.Pathtracer(arrays, cfg).Init()
pt.Render(cams[0], shouldClear=True)
for i in range(5): pt.Render(cams[i % 16])
pt.Finish(); pt.set_profiling(True)
for i in range(50): pt.Render(cams[i % 16])
pt.Finish()
print("  training stage us/frame:", {k: round(1e3 * v[0] / 50, 1) for k, v in pt.profile().items()}, "launches/frame", sum(v[1] for v in pt.profile().values()) / 50)
pt.close()
