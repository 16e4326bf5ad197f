#!/bin/bash
# The GPU suite on the bounds-checked build of the library (stands in for compute-sanitizer memcheck, which the pool does not offer).
# Build here first: make -C cuda_pathtracer_b200/csrc variant NAME=_chk DEFS=-DPT_BOUNDS_CHECK
mkdir -p gpurun_out
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_chk.so
timeout 420 python -m pytest tests -x -q -m gpu --timeout 120 > gpurun_out/pytest_gpu_bounds_checked.log 2>&1; echo "checked suite rc=$?"; tail -n 4 gpurun_out/pytest_gpu_bounds_checked.log
PTB200_COOP=100000 timeout 240 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_round2.py tests/test_zy_bounds_checked.py -x -q -m gpu --timeout 120 > gpurun_out/pytest_gpu_bounds_checked_coop.log 2>&1; echo "checked, everything cooperative rc=$?"; tail -n 3 gpurun_out/pytest_gpu_bounds_checked_coop.log
