#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_bvh_build.py tests/test_gpu_trace.py tests/test_gpu_baseline_configs.py tests/test_gpu_edge_cases.py tests/test_gpu_render.py -x -q -m gpu --timeout 200 > gpurun_out/r2dp2_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2dp2_pytest.log
AB_TALLY=1 AB_WORLDS=1,8 timeout 500 python tools/ab_probe.py "greedy=PTB200_WIDE_GREEDY=1" "optimal=" "greedy2=PTB200_WIDE_GREEDY=1" "optimal2=" > gpurun_out/r2dp2_ab.log 2>&1
cut -c1-300 gpurun_out/r2dp2_ab.log
python tools/bvh_build_bench.py > gpurun_out/r2dp2_build.log 2>&1; tail -n 5 gpurun_out/r2dp2_build.log | cut -c1-400
