#!/bin/bash
# bench per environment setting. usage: gpurun -- 'bash tools/tune_env.sh tag "A=1" "B=2 C=3" ...'  ("-" = no setting)
TAG=$1; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for e in "$@"; do
  echo "== env $e"
  [ "$e" = "-" ] && e=""
  env $e python bench.py --steps 100 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(json.dumps({k:d[k] for k in ('value','ms_per_step','stage_ms_per_step','gpu_launches')})); print(d['e2e']['value'], d['roofline']['frac'])
    else: print(l.rstrip()[:300])
" | tee -a gpurun_out/${TAG}_tune.log
done
