#!/bin/bash
# usage: gpurun -- 'bash tools/tune2.sh tag _t1 _t3 ...' : bench line (value / training / steady) per library variant, then the
# hit-id parity tests with every variant.
TAG=$1; shift
mkdir -p gpurun_out
for v in "" "$@"; do
  echo "== variant '$v'" | tee -a gpurun_out/${TAG}_tune.log
  PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so python bench.py --steps 100 --warmup 5 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1f  training %.4f ms  steady %.4f ms  e2e %.1f' % (d['value'], d['phases']['training_ms_per_step'], d['phases']['steady_ms_per_step'], d['e2e']['value'])); print(json.dumps(d['stage_ms_per_step']))
" | tee -a gpurun_out/${TAG}_tune.log
done
for v in "$@"; do
  echo "== parity with variant '$v'" | tee -a gpurun_out/${TAG}_tune.log
  PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -x -q 2>&1 | tail -2 | tee -a gpurun_out/${TAG}_tune.log
done
