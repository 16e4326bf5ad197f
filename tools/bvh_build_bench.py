#!/usr/bin/env python
"""Host vs device BVH2 build time (same tree): tools/bvh_build_bench.py [out.json].  The reference quotes 686 ms for
2.0 M triangles with its OpenMP host builder (screenshots/2mtris.png)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from cuda_pathtracer_b200 import api, scenes
from cuda_pathtracer_b200.scene import Scene, material
rows = []
for name, make in (("statue 20k", scenes.statue_soup), ("cathedral 138k", scenes.cathedral_soup), ("heightfield 2.0M (2mtris proxy)", scenes.heightfield_soup)):
    v = np.ascontiguousarray(make(), np.float32).reshape(-1, 9)
    api.build_bvh(v[:1000])                                   # context + module load outside the timing
    best = None
    for _ in range(3):
        t0 = time.perf_counter(); nodes, order, ms, levels = api.build_bvh(v); wall = (time.perf_counter() - t0) * 1e3
        best = (ms, wall, levels, len(nodes)) if best is None or ms < best[0] else best
    s = Scene(); s.addMaterial(material()); t0 = time.perf_counter(); s.addModelSoup(v); host_wall = (time.perf_counter() - t0) * 1e3
    host_ms = s.last_build_ms()
    s.addObject(0); s.finalize()
    arrays = s.arrays()
    same = len(arrays["bvh"][0]) == best[3]
    from cuda_pathtracer_b200 import abi
    up = {}
    for label, flags in (("warm-up", 0), ("device_wide_bvh", 0), ("host_wide_bvh", abi.FLAG_HOST_WIDE_BVH)):
        pt = api.Pathtracer(arrays, abi.default_config(64, 64, max_ray_depth=2, interactive_depth=1, flags=flags))
        t0 = time.perf_counter(); pt.Init(); up[label] = (time.perf_counter() - t0) * 1e3
        pt.close()
    rows.append({"mesh": name, "triangles": len(v), "nodes": best[3], "levels": best[2], "device_build_ms": round(best[0], 3), "device_call_wall_ms_incl_copies_and_alloc": round(best[1], 2),
                 "host_build_ms": round(host_ms, 1), "speedup_vs_host_builder": round(host_ms / best[0], 1), "same_node_count": bool(same),
                 "upload_scene_wall_ms_device_wide_bvh": round(up["device_wide_bvh"], 1), "upload_scene_wall_ms_host_wide_bvh": round(up["host_wide_bvh"], 1)})
    print(rows[-1], flush=True)
if len(sys.argv) > 1:
    json.dump({"what": "binned-SAH BVH2 build, device (pt_build_bvh) vs host builder (host_scene.cpp, 1 thread), and pt_upload_scene with the BVH2 -> 8-wide conversion on the device vs on the host; identical outputs (tests/test_gpu_bvh_build.py)", "rows": rows}, open(sys.argv[1], "w"), indent=1)
