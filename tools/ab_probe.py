#!/usr/bin/env python
"""A/B of scheduling knobs on ONE GPU: training ms/frame and steady ms/sample of the bench workload (sibenik 1080p, depth 8)
for the whole image (world 1) and for rank 0's share of an emulated 8-rank job, one line per variant.
usage: tools/ab_probe.py [name=ENV1=v;ENV2=v ...]   (default: a fixed list)"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cuda_pathtracer_b200 import abi, api, scenes

VARIANTS = [("base", {}), ("no_pdl", {"PTB200_NO_PDL": "1"}), ("sort4x2", {"PTB200_SORT": "4,2"}), ("sort4x1", {"PTB200_SORT": "4,1"}),
            ("sort5x2", {"PTB200_SORT": "5,2"}), ("sort3x2", {"PTB200_SORT": "3,2"})]
if len(sys.argv) > 1:
    VARIANTS = []
    for a in sys.argv[1:]:
        name, _, rest = a.partition("=")
        VARIANTS.append((name, dict(kv.split("=", 1) for kv in rest.split(";") if kv) if rest else {}))
arrays = scenes.sibenik_scene().arrays()
cams = scenes.camera_sweep("sibenik", 1920, 1080, 16)
KNOBS = ["PTB200_NO_PDL", "PTB200_SORT", "PTB200_TAIL_BOUNCE", "PTB200_MIN_WIDTH", "PTB200_SETS", "PTB200_NO_TAIL_PAIR", "PTB200_COOP", "PTB200_TAIL_GRID", "PTB200_TAIL_PAIR_MAX", "PTB200_PDL_ALWAYS", "PTB200_NO_MIX", "PTB200_TAIL_ALONE", "PTB200_NO_FEEDBACK", "PTB200_HEAD_AT", "PTB200_ANY_FAR_FIRST", "PTB200_SPLIT", "PTB200_HOST_WIDE_BVH", "PTB200_WIDE_GREEDY", "PTB200_LEAF_INDEX_CUT"]


def measure(world, env):
    for k in KNOBS:
        os.environ.pop(k, None)
    os.environ.update(env)
    cfg = abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1, rank=0, world_size=world)
    pt = api.Pathtracer(arrays, cfg).Init()
    stream = torch.cuda.ExternalStream(pt.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pt.set_guiding_frames(1 << 30)
    pt.Render(cams[0], shouldClear=True)
    for i in range(5): pt.Render(cams[i % 16])
    best_t = 1e9
    for rep in range(3):
        pt.join(); e0.record(stream)
        for i in range(40): pt.Render(cams[i % 16])
        pt.join(); e1.record(stream); pt.Finish()
        best_t = min(best_t, e0.elapsed_time(e1) / 40)
    stages = None
    if os.environ.get("AB_STAGES"):
        pt.set_profiling(True)
        for i in range(20): pt.Render(cams[i % 16])
        pt.Finish()
        stages = {k: round(1e3 * v[0] / 20, 1) for k, v in pt.profile().items()}
        pt.set_profiling(False)
    pt.set_guiding_frames(0)
    for i in range(16): pt.Render(cams[i % 16])
    pt.Finish(); pt.reset_stats()
    best_s, all_s = 1e9, []
    for rep in range(5):
        pt.join(); e0.record(stream)
        for i in range(64): pt.Render(cams[i % 16])
        pt.join(); e1.record(stream); pt.Finish()
        all_s.append(round(e0.elapsed_time(e1) / 64, 4))
        best_s = min(best_s, all_s[-1])
    st = pt.stats()
    rays = (st.extend_rays + st.shadow_rays) / (5 * 64)
    if os.environ.get("AB_TALLY"):            # work per ray as the counting kernels see it
        pt.Finish(); pt.reset_stats(); pt.set_counting(True)
        for i in range(4): pt.Render(cams[i % 16])
        pt.Finish()
        t, s2 = pt.tally(), pt.stats()
        stages = dict(stages or {}, tally_per_extend_ray=[round(float(x) / max(1, s2.extend_rays), 3) for x in t["extend"]],
                      tally_per_shadow_ray=[round(float(x) / max(1, s2.shadow_rays), 3) for x in t["connect"]])
        pt.set_counting(False)
    pt.close()
    return best_t, best_s, rays, stages, all_s


WORLDS = [int(w) for w in os.environ.get("AB_WORLDS", "1,8").split(",")]
for world in WORLDS:
    for name, env in VARIANTS:
        t, s, rays, stages, all_s = measure(world, env)
        print(json.dumps({"world": world, "variant": name, "training_ms": round(t, 4), "steady_ms": round(s, 4), "rays_per_sample": int(rays),
                          "training_mrays_s": round(rays / t / 1e3, 1), "steady_mrays_s": round(rays / s / 1e3, 1), "steady_windows": all_s,
                          "training_stage_us": stages}), flush=True)
