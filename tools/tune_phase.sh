#!/bin/bash
# bench phases per environment setting. usage: gpurun -- 'bash tools/tune_phase.sh tag "A=1" "B=2 C=3" ...'  ("-" = no setting)
TAG=$1; shift
mkdir -p gpurun_out
for e in "$@"; do
  echo "== env $e" | tee -a gpurun_out/${TAG}_phase.log
  [ "$e" = "-" ] && e=""
  env $e python bench.py --steps 120 --warmup 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); p=d['phases']; print('value %.0f  training %.0f (%.3f ms)  steady %.0f (%.3f ms)  e2e %.0f' % (d['value'], p['training_mrays_s'], p['training_ms_per_step'], p['steady_mrays_s'], p['steady_ms_per_step'], d['e2e']['value']))
    else: print(l.rstrip()[:200])
" | tee -a gpurun_out/${TAG}_phase.log
done
