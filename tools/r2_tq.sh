#!/bin/bash
# warp-wide triangle queue (build knob PT_TRI_QUEUE): parity, then A/B against the default library
TAG=${1:-r2tq}
mkdir -p gpurun_out
AB_STAGES=1 AB_WORLDS=1,8 timeout 600 python tools/ab_probe.py 'base=' > gpurun_out/${TAG}_ab_base.log 2>&1; cut -c1-500 gpurun_out/${TAG}_ab_base.log
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_tq.so
timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_vs_reference_device.py tests/test_gpu_edge_cases.py tests/test_gpu_baseline_configs.py -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/${TAG}_pytest.log
AB_STAGES=1 AB_WORLDS=1,8 timeout 600 python tools/ab_probe.py 'tq=' > gpurun_out/${TAG}_ab_tq.log 2>&1; cut -c1-500 gpurun_out/${TAG}_ab_tq.log
