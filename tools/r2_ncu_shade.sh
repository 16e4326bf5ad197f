#!/bin/bash
# full ncu capture of the shade / primary / accumulate / cache launches of one frame of the bench command
TAG=${1:-r2s}
mkdir -p gpurun_out
timeout 100 python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 500 ncu --set full --clock-control none --import-source on -k "regex:k_shade|k_primary|k_accumulate|k_rc_" -s 10 -c 14 -o gpurun_out/${TAG}_shade -f python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out | grep ${TAG}
