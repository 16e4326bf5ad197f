#!/bin/bash
# Runs the bench for each tuning variant of the library (built beforehand with `make variant`).
# usage: gpurun -- 'bash tools/tune.sh tag _t1 _t2 ...'
TAG=$1; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in "" "$@"; do
  echo "== variant '$v'"
  PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so python bench.py --steps 60 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(json.dumps({k:d[k] for k in ('value','ms_per_step','stage_ms_per_step')})); print('steady', d['phases']['steady_mrays_s'], 'e2e', d['e2e']['value'])
    else: print(l.rstrip()[:300])
" | tee -a gpurun_out/${TAG}_tune.log
done
