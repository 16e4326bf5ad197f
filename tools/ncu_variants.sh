#!/bin/bash
# Instruction counts / lane utilisation of the traversal launches of one sample, per tuning variant.
# usage: gpurun -- 'bash tools/ncu_variants.sh tag "" _v1 _v2'
TAG=$1; shift
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 || { echo plain run failed; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
for v in "$@"; do
  PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200$v.so ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active \
     --clock-control none -k regex:k_trace -s 50 -c 16 --csv --log-file gpurun_out/${TAG}_ncu$v.csv python bench.py --steps 2 --warmup 3 > /dev/null 2>&1
  echo "variant '$v' ncu exit $?"
done
