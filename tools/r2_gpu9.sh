#!/bin/bash
TAG=${1:-r2n}
mkdir -p gpurun_out
S=$(date +%s)
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_far.so
PTB200_ANY_FAR_FIRST=1 timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_vs_reference_device.py tests/test_gpu_edge_cases.py -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -3 gpurun_out/${TAG}_pytest.log
AB_STAGES=1 AB_WORLDS=1,8 timeout 900 python tools/ab_probe.py 'near=' 'far=PTB200_ANY_FAR_FIRST=1' 'near2=' 'far2=PTB200_ANY_FAR_FIRST=1' > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cut -c1-420 gpurun_out/${TAG}_ab.log
