#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" _cpf; do
  export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so
  AB_STAGES=1 AB_WORLDS=8,4,1 timeout 600 python tools/ab_probe.py "lib$v=" >> gpurun_out/r2cpf_ab.log 2>&1
done
done
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_cpf.so
PTB200_COOP=100000 timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -x -q -m gpu > gpurun_out/r2cpf_pytest.log 2>&1; echo "pytest rc=$?"
cut -c1-420 gpurun_out/r2cpf_ab.log
