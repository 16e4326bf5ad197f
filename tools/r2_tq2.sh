#!/bin/bash
mkdir -p gpurun_out
for v in "" _tq _tq16 _tq8 _tqp1 _tqp6; do
  export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so
  AB_TALLY=1 AB_WORLDS=1 timeout 600 python tools/ab_probe.py "lib$v=" >> gpurun_out/r2tq2_ab.log 2>&1
done
cut -c1-700 gpurun_out/r2tq2_ab.log
