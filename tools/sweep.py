#!/usr/bin/env python
"""Camera-sweep throughput of the BASELINE.json configs on one GPU: K cameras per scene, per camera
a clearing frame + warm-up samples + timed converge samples (CUDA events on the context's stream), reporting min /
median / max Mrays/s and samples/s over the cameras.  usage: tools/sweep.py [config ...] [--cams K] [--samples S]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from cuda_pathtracer_b200 import abi, api, scenes

CONFIGS = {
    # name: (scene factory, camera set, width, height, depth, nee, guiding, textures?)
    "C2_outside_1080p": dict(scene=lambda: scenes.outside_scene(), cams="outside", w=1920, h=1080, depth=8, nee=True, guiding=False, interactive=5),
    "C3_sibenik_1080p": dict(scene=lambda: scenes.sibenik_scene(), cams="sibenik", w=1920, h=1080, depth=8, nee=True, guiding=True),
    "C4_2mtris_4k": dict(scene=lambda: scenes.two_million_scene(), cams="2mtris", w=3840, h=2160, depth=8, nee=False, guiding=False),
    "C5_minecraft_1080p": dict(scene=lambda: scenes.minecraft_scene(256, copies=20), cams="minecraft", w=1920, h=1080, depth=8, nee=True, guiding=False, textured=True),
}

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=list(CONFIGS))
    ap.add_argument("--cams", type=int, default=16)
    ap.add_argument("--warm", type=int, default=8)
    ap.add_argument("--samples", type=int, default=32)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rows = []
    for name in args.configs:
        c = CONFIGS[name]
        t0 = time.time(); s = c["scene"](); arrays = s.arrays(); build_s = time.time() - t0
        cfg = abi.default_config(c["w"], c["h"], max_ray_depth=c["depth"], interactive_depth=c.get("interactive", 1))
        pt = api.Pathtracer(arrays, cfg)
        t0 = time.time()
        pt.Init(textures={s.textured_material: scenes.voxel_atlas(256)} if c.get("textured") else None)
        upload_s = time.time() - t0
        if c.get("textured"):
            pt.set_blue_noise(np.random.default_rng(7).random((256, 256)).astype(np.float32))
        pt.HNEE, pt.HCACHE = c["nee"], c["guiding"]
        stream = torch.cuda.ExternalStream(pt.stream())
        if c["cams"] == "minecraft":
            base = scenes.minecraft_camera(c["w"], c["h"], 256)
            cams = [base] * 1 + [scenes.make_camera((0.35 * 256 * np.cos(a) + 10 * 128, 14.0 + 4 * np.sin(3 * a), -0.35 * 256 * np.sin(a)),
                                                    tuple(np.array([-np.cos(a), -0.25, np.sin(a)]) / np.linalg.norm([np.cos(a), 0.25, np.sin(a)])), 1.5, 90.0, 0.05, c["w"], c["h"])
                                 for a in np.linspace(0.3, 2.8, args.cams - 1)]
        else:
            cams = scenes.camera_sweep(c["cams"], c["w"], c["h"], args.cams)
        if c["guiding"]:                      # train the cache like the reference does (first 100 converge frames) before the sweep
            pt.Render(cams[0], shouldClear=True)
            for i in range(100): pt.Render(cams[i % len(cams)])
            pt.Finish(); pt.set_guiding_frames(0)
        mr, sps, rpp = [], [], []
        for cam in cams:
            pt.Render(cam, shouldClear=True)
            for _ in range(args.warm): pt.Render(cam)
            pt.Finish(); pt.reset_stats()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.samples): pt.Render(cam)
            pt.join(); e1.record(stream); pt.Finish()
            ms = e0.elapsed_time(e1); st = pt.stats()
            rays = st.extend_rays + st.shadow_rays
            mr.append(rays / ms / 1e3); sps.append(args.samples / ms * 1e3); rpp.append(rays / args.samples / (c["w"] * c["h"]))
        img = pt.read_image()
        row = {"config": name, "triangles": int(len(arrays["vertices"])), "instances": int(len(arrays["instances"])), "pixels": c["w"] * c["h"],
               "scene_build_s": round(build_s, 2), "upload_s": round(upload_s, 2), "cameras": len(cams), "samples_per_camera": args.samples,
               "mrays_s": {"min": min(mr), "median": float(np.median(mr)), "max": max(mr)},
               "samples_s": {"min": min(sps), "median": float(np.median(sps)), "max": max(sps)},
               "rays_per_pixel_sample": float(np.mean(rpp)), "finite": bool(np.isfinite(img).all()), "w_ok": bool(np.all(img[..., 3] == args.samples + args.warm + cfg.interactive_depth))}
        print(json.dumps(row), flush=True)
        rows.append(row)
        pt.close()
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)

if __name__ == "__main__":
    main()
