#!/bin/bash
TAG=${1:-r2c}
mkdir -p gpurun_out
S=$(date +%s)
timeout 600 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_round2.py -q -s > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; grep -E "C4|ray [0-9]+:|rel RMSE|passed|failed|mean radiance|sky sampling" gpurun_out/${TAG}_pytest.log | head -30
AB_STAGES=1 AB_WORLDS=1,2,4,8 timeout 900 python tools/ab_probe.py base= t8=PTB200_COOP=8,1 t16=PTB200_COOP=16,1 t32=PTB200_COOP=32,1 t16x2=PTB200_COOP=16,2 t16x4=PTB200_COOP=16,4 t64x2=PTB200_COOP=64,2 t16wave=PTB200_COOP=16,1,PTB200_TAIL_BOUNCE=8 no_pdl=PTB200_NO_PDL=1 > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cat gpurun_out/${TAG}_ab.log
