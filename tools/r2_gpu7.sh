#!/bin/bash
TAG=${1:-r2h}
mkdir -p gpurun_out
S=$(date +%s)
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_vs_reference_device.py tests/test_gpu_edge_cases.py tests/test_gpu_peer_exchange.py -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -3 gpurun_out/${TAG}_pytest.log
AB_STAGES=1 AB_WORLDS=1,2,4,8 timeout 900 python tools/ab_probe.py 'feedback=' 'fixed3=PTB200_NO_FEEDBACK=1' > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cut -c1-420 gpurun_out/${TAG}_ab.log
