#!/bin/bash
mkdir -p gpurun_out
PTB200_HOST_WIDE_BVH=1 PTB200_WIDE_DP=0.3 timeout 300 python -m pytest tests/test_gpu_trace.py tests/test_gpu_baseline_configs.py -x -q -m gpu --timeout 120 > gpurun_out/r2dp_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2dp_pytest.log
AB_TALLY=1 AB_WORLDS=1 timeout 500 python tools/ab_probe.py "greedy=PTB200_HOST_WIDE_BVH=1" "dp0.3=PTB200_HOST_WIDE_BVH=1;PTB200_WIDE_DP=0.3" "dp0.15=PTB200_HOST_WIDE_BVH=1;PTB200_WIDE_DP=0.15" "dp0.6=PTB200_HOST_WIDE_BVH=1;PTB200_WIDE_DP=0.6" "dp1.0=PTB200_HOST_WIDE_BVH=1;PTB200_WIDE_DP=1.0" > gpurun_out/r2dp_ab.log 2>&1
cut -c1-600 gpurun_out/r2dp_ab.log
