#!/usr/bin/env python
"""torchrun --nproc-per-node N tools/multi_gpu_check.py : N tile-sharded ranks with the NCCL all-reduce hook against the
same render on one rank (computed on rank 0's GPU).  Guiding off -> the gathered framebuffer must be BIT-IDENTICAL;
guiding on (60 training frames, float atomics + a different summation order) -> relative RMSE bound + caches replicated."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
from cuda_pathtracer_b200 import abi, api, scenes
from cuda_pathtracer_b200.distributed import make_allreduce_hook, gather_image

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
W, H = 480, 270
arrays = scenes.sibenik_scene().arrays()
cam = scenes.camera_sweep("sibenik", W, H, 16)[2]
out = {}
for guiding, frames in ((False, 8), (True, 60)):
    def render(r, w, hook):
        cfg = abi.default_config(W, H, max_ray_depth=8, interactive_depth=1, device=local, rank=r, world_size=w)
        pt = api.Pathtracer(arrays, cfg).Init()
        if hook:
            pt.set_allreduce_hook(make_allreduce_hook(torch.device("cuda", local)))
        pt.HNEE, pt.HCACHE = True, guiding
        pt.Render(cam, shouldClear=True)
        for _ in range(frames):
            pt.Render(cam)
        pt.Finish()
        img = pt.read_image().copy(); rc = pt.read_radiance_cache()["radianceCache"].copy()
        pt.close()
        return img, rc
    img, rc = render(rank, world, True)
    full = gather_image(torch.from_numpy(img).cuda()).cpu().numpy()
    rcs = [torch.empty_like(torch.from_numpy(rc).cuda()) for _ in range(world)]
    dist.all_gather(rcs, torch.from_numpy(rc).cuda())
    if rank == 0:
        ref, rc1 = render(0, 1, False)
        a, b = full[..., :3] / full[..., 3:4], ref[..., :3] / ref[..., 3:4]
        err = float(np.sqrt(np.mean((a.astype(np.float64) - b) ** 2)) / np.sqrt(np.mean(b.astype(np.float64) ** 2)))
        replicated = all(bool(torch.equal(rcs[0], x)) for x in rcs[1:])
        out["guiding_on" if guiding else "guiding_off"] = {"frames": frames, "bit_identical": bool(np.array_equal(full, ref)), "rel_rmse": err,
                                                           "samples_ok": bool(np.all(full[..., 3] == frames + 1)), "caches_replicated_bitwise": replicated,
                                                           "cache_rel_rmse_vs_1rank": float(np.sqrt(np.mean((rcs[0].cpu().numpy() - rc1) ** 2)) / np.sqrt(np.mean(rc1 ** 2)))}
if rank == 0:
    ok = out["guiding_off"]["bit_identical"] and out["guiding_on"]["rel_rmse"] < 2e-2 and out["guiding_on"]["samples_ok"]
    print(json.dumps({"world": world, "ok": bool(ok), **out}))
dist.destroy_process_group()
