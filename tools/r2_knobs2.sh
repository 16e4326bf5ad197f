#!/bin/bash
mkdir -p gpurun_out
for v in "" _mb9 _mb10 _mb6 _sh6 _sh10 ""; do
  export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so
  AB_WORLDS=1,8 timeout 300 python tools/ab_probe.py "lib$v=" >> gpurun_out/r2knobs2_ab.log 2>&1
done
cut -c1-160 gpurun_out/r2knobs2_ab.log
