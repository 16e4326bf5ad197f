#!/bin/bash
TAG=${1:-r2e}
mkdir -p gpurun_out
S=$(date +%s)
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_gpu_trace.py tests/test_gpu_render.py -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -3 gpurun_out/${TAG}_pytest.log
AB_WORLDS=1,2,4,8 timeout 900 python tools/ab_probe.py 'default=' 'off=PTB200_COOP=0' 'c4x1=PTB200_COOP=4,1' 'c4x4=PTB200_COOP=4,4' 'c8x4=PTB200_COOP=8,4' 'c2x2=PTB200_COOP=2,2' > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cat gpurun_out/${TAG}_ab.log | cut -c1-330
for v in 0 "4,2" "8,8"; do
  PTB200_COOP=$v timeout 100 python tools/shard_launches.py 8 > gpurun_out/${TAG}_shard_plain_$v.log 2>&1 &&
  PTB200_COOP=$v timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_shard8_coop$v.csv python tools/shard_launches.py 8 > /dev/null 2>&1
  echo "ncu shard coop=$v exit $? $(( $(date +%s)-S ))s"; tail -1 gpurun_out/${TAG}_shard_plain_$v.log | cut -c1-400
done
