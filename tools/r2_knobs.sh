#!/bin/bash
mkdir -p gpurun_out
for v in "" _t1 _t3 _t4 _rf2 _rf4 ""; do
  export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so
  AB_WORLDS=1,8 timeout 300 python tools/ab_probe.py "lib$v=" >> gpurun_out/r2knobs_ab.log 2>&1
done
cut -c1-160 gpurun_out/r2knobs_ab.log
