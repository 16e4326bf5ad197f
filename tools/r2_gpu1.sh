#!/bin/bash
# First GPU session of round 2: parity suite, bench line as the driver runs it, knob A/B, per-bounce queue sizes.
TAG=${1:-r2a}
mkdir -p gpurun_out
S=$(date +%s)
timeout 900 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -25 gpurun_out/${TAG}_pytest.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? $(( $(date +%s)-S ))s"; cut -c1-600 gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
timeout 300 python tools/ab_probe.py > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cat gpurun_out/${TAG}_ab.log
timeout 100 python - <<'P' 2>&1 | tail -5
import sys; sys.path.insert(0, '.')
from cuda_pathtracer_b200 import abi, api, scenes
arrays = scenes.sibenik_scene().arrays(); cams = scenes.camera_sweep("sibenik", 1920, 1080, 16)
for world in (1, 8):
    pt = api.Pathtracer(arrays, abi.default_config(1920, 1080, max_ray_depth=8, interactive_depth=1, rank=0, world_size=world)).Init()
    pt.Render(cams[0], shouldClear=True); pt.set_counting(True)
    for i in range(3): pt.Render(cams[i])
    pt.Finish(); e, s = pt.debug_read_counters(8); print("world", world, "ext", list(e), "shadow", list(s), "tally", pt.tally()); pt.close()
P
