#!/bin/bash
# split cache-training frames: parity under PTB200_SPLIT, then the A/B
mkdir -p gpurun_out
PTB200_SPLIT=4 timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_vs_reference_device.py tests/test_gpu_round2.py -x -q -m gpu > gpurun_out/split_pytest.log 2>&1
echo "pytest split rc=$?"; tail -n 5 gpurun_out/split_pytest.log
AB_WORLDS=1,8 timeout 1200 python tools/ab_probe.py base= "s2=PTB200_SPLIT=2" "s4=PTB200_SPLIT=4" "s4w1=PTB200_SPLIT=4,1" "s4notail=PTB200_SPLIT=4,32,0" "s4hold=PTB200_SPLIT=4,32,1,1" "s2w1=PTB200_SPLIT=2,1" "s3=PTB200_SPLIT=3" > gpurun_out/ab_split.log 2>&1
cat gpurun_out/ab_split.log | cut -c1-200
