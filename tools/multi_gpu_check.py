#!/usr/bin/env python
"""torchrun --nproc-per-node N tools/multi_gpu_check.py [--same-gpu] [--frames F] [--size WxH]
N tile-sharded ranks against the same render on one rank (computed on rank 0's GPU), once with the all-reduce hook and once
with the fused reduce + propagate kernel over peer memory (connect_peers).  Guiding off -> the gathered framebuffer must be
BIT-IDENTICAL; guiding on (float atomics + a different summation order) -> relative RMSE bound + caches replicated bitwise.
--same-gpu puts every rank on device 0 (gloo carries the handles and the hook): the peer protocol -- cudaIpc mapping,
flags, owner ranges -- then runs between two PROCESSES of one GPU, which is how the 1-GPU test box covers it
(tests/test_gpu_peer_exchange.py); across GPUs the same kernels run over NVLink."""
import argparse, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
from cuda_pathtracer_b200 import abi, api, scenes
from cuda_pathtracer_b200.distributed import make_allreduce_hook, connect_peers, disconnect_peers, HostRing

ap = argparse.ArgumentParser()
ap.add_argument("--same-gpu", action="store_true")
ap.add_argument("--frames", type=int, default=60)
ap.add_argument("--size", default="480x270")
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
if args.same_gpu:
    local = 0
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if args.same_gpu:
    dist.init_process_group("gloo")
else:
    dist.init_process_group("nccl", device_id=dev)
W, H = (int(v) for v in args.size.split("x"))
arrays = scenes.sibenik_scene().arrays()
cam = scenes.camera_sweep("sibenik", W, H, 16)[2]


def gather_sum(a):      # every rank holds zeros outside its tiles: the sum is the framebuffer
    t = torch.from_numpy(a) if args.same_gpu else torch.from_numpy(a).to(dev)
    dist.all_reduce(t)
    return t.cpu().numpy()


def gather_all(a):
    parts = [None] * world
    dist.all_gather_object(parts, a)
    return parts


out = {}
for guiding, frames, transport in ((False, 8, "hook"), (True, args.frames, "hook"), (True, args.frames, "peer")):
    def render(r, w, multi):
        cfg = abi.default_config(W, H, max_ray_depth=8, interactive_depth=1, device=local, rank=r, world_size=w)
        pt = api.Pathtracer(arrays, cfg).Init()
        if multi and transport == "hook":
            pt.set_allreduce_hook(make_allreduce_hook(dev))
        elif multi:
            assert connect_peers(pt), "peer mapping failed"
        pt.HNEE, pt.HCACHE = True, guiding
        pt.Render(cam, shouldClear=True)
        for _ in range(frames):
            pt.Render(cam)
        gathered = None
        if multi and transport == "peer":              # one framebuffer on rank 0: tiles over NVLink into its image ring, one copy to the host
            host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True).numpy() if r == 0 else None
            for _ in range(6):                          # more epochs than ring slots: exercises the slot-reuse handshake
                t = pt.gather_image_async(host)
            pt.wait_read(t)
            gathered = host.copy() if r == 0 else None
            # the same framebuffer through the shared host ring: every rank writes its tiles through its own PCIe link
            hr = HostRing(pt, slots=3)
            for k in range(7):                          # more epochs than slots: the consumer's release gates slot reuse
                e = pt.deliver_image_async()
                if r == 0:
                    view = pt.host_ring_wait(e)
                    delivered = view.copy()
                    pt.host_ring_release(e)
            gathered = (gathered, delivered) if r == 0 else None
            hr.close()
        pt.Finish()
        img = pt.read_image().copy(); rc = pt.read_radiance_cache()["radianceCache"].copy()
        if multi and transport == "peer":
            disconnect_peers(pt)
        pt.close()
        return (img, rc, gathered) if multi else (img, rc)
    img, rc, gathered = render(rank, world, True)
    full = gather_sum(img)
    rcs = gather_all(rc)
    if rank == 0:
        ref, rc1 = render(0, 1, False)
        a, b = full[..., :3] / full[..., 3:4], ref[..., :3] / ref[..., 3:4]
        err = float(np.sqrt(np.mean((a.astype(np.float64) - b) ** 2)) / np.sqrt(np.mean(b.astype(np.float64) ** 2)))
        replicated = all(bool(np.array_equal(rcs[0], x)) for x in rcs[1:])
        out[("guiding_on_" + transport) if guiding else "guiding_off"] = {
            "frames": frames, "bit_identical": bool(np.array_equal(full, ref)), "rel_rmse": err,
            "samples_ok": bool(np.all(full[..., 3] == frames + 1)), "caches_replicated_bitwise": replicated,
            "cache_moved": bool(np.any(rc1 != 0.1)),
            "gathered_image_identical": None if gathered is None else bool(np.array_equal(gathered[0], full)),
            "host_ring_image_identical": None if gathered is None else bool(np.array_equal(gathered[1], full)),
            "cache_rel_rmse_vs_1rank": float(np.sqrt(np.mean((rcs[0] - rc1) ** 2)) / np.sqrt(np.mean(rc1 ** 2)))}
    dist.barrier()
if rank == 0:
    ok = out["guiding_off"]["bit_identical"] and out["guiding_on_peer"]["gathered_image_identical"] and out["guiding_on_peer"]["host_ring_image_identical"] and all(out[k]["rel_rmse"] < 2e-2 and out[k]["samples_ok"] and out[k]["caches_replicated_bitwise"] and out[k]["cache_moved"]
                                                     for k in ("guiding_on_hook", "guiding_on_peer"))
    print(json.dumps({"world": world, "same_gpu": args.same_gpu, "ok": bool(ok), **out}))
dist.destroy_process_group()
