#!/bin/bash
TAG=${1:-r2i}
mkdir -p gpurun_out
S=$(date +%s)
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_gpu_render.py tests/test_gpu_vs_reference_device.py -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -3 gpurun_out/${TAG}_pytest.log
AB_STAGES=1 AB_WORLDS=1,2,4,8 timeout 900 python tools/ab_probe.py 'feedback=' 'head2=PTB200_HEAD_AT=2' 'head4=PTB200_HEAD_AT=4' 'head8=PTB200_HEAD_AT=8' 'fixed3=PTB200_NO_FEEDBACK=1' > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cut -c1-420 gpurun_out/${TAG}_ab.log
