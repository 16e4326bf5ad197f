#!/bin/bash
# bench per tail-bounce setting (PTB200_TAIL_BOUNCE developer knob). usage: gpurun -- 'bash tools/tune_tail.sh tag 64 2 3 4'
TAG=$1; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for t in "$@"; do
  echo "== tail bounce $t"
  PTB200_TAIL_BOUNCE=$t python bench.py --steps 60 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(json.dumps({k:d[k] for k in ('value','ms_per_step','stage_ms_per_step','gpu_launches')})); print(d['e2e']['value'], d['roofline']['frac'])
    else: print(l.rstrip()[:300])
" | tee -a gpurun_out/${TAG}_tune.log
done
