#!/bin/bash
TAG=${1:-r2g}
mkdir -p gpurun_out
S=$(date +%s)
AB_STAGES=1 AB_WORLDS=1,2,4,8 timeout 900 python tools/ab_probe.py 'tail3=' 'tail2=PTB200_TAIL_ALONE=2' 'tail4=PTB200_TAIL_ALONE=4' 'tail5=PTB200_TAIL_ALONE=5' 'tail6=PTB200_TAIL_ALONE=6' 'wave=PTB200_TAIL_ALONE=8' > gpurun_out/${TAG}_ab.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cut -c1-420 gpurun_out/${TAG}_ab.log
