#!/bin/bash
TAG=${1:-r2c}
mkdir -p gpurun_out
S=$(date +%s)
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_next.so
AB_STAGES=1 AB_WORLDS=1,8 timeout 900 python tools/ab_probe.py 'base=' 't8=PTB200_COOP=8,1' 't16=PTB200_COOP=16,1' 't32=PTB200_COOP=32,1' 't16x2=PTB200_COOP=16,2' 't16x4=PTB200_COOP=16,4' 't64x2=PTB200_COOP=64,2' 't16wave=PTB200_COOP=16,1;PTB200_TAIL_BOUNCE=8' 'pair24=PTB200_TAIL_PAIR_MAX=24' 'pair32=PTB200_TAIL_PAIR_MAX=32' 'pair48=PTB200_TAIL_PAIR_MAX=48' 'tail4=PTB200_TAIL_BOUNCE=4' 'tail4c=PTB200_TAIL_BOUNCE=4;PTB200_COOP=16,2' 'tail2=PTB200_TAIL_BOUNCE=2' > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cat gpurun_out/${TAG}_ab.log
AB_WORLDS=2,4 timeout 900 python tools/ab_probe.py 'base=' 't16x2=PTB200_COOP=16,2' 't32x4=PTB200_COOP=32,4' > gpurun_out/${TAG}_ab24.log 2>&1; cat gpurun_out/${TAG}_ab24.log
