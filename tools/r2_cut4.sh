#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_bvh_build.py tests/test_gpu_trace.py tests/test_gpu_baseline_configs.py tests/test_gpu_edge_cases.py tests/test_gpu_render.py tests/test_gpu_round2.py -x -q -m gpu --timeout 200 > gpurun_out/r2cut4_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2cut4_pytest.log
AB_TALLY=1 AB_WORLDS=1,8 timeout 500 python tools/ab_probe.py "index_cut=PTB200_LEAF_INDEX_CUT=1" "best_cut=" "index_cut2=PTB200_LEAF_INDEX_CUT=1" "best_cut2=" > gpurun_out/r2cut4_ab.log 2>&1
cut -c1-200 gpurun_out/r2cut4_ab.log
