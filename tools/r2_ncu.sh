#!/bin/bash
# full ncu capture of one whole frame's traversal launches of the bench command (after the same command ran clean)
TAG=${1:-r2p}
mkdir -p gpurun_out
timeout 100 python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 500 ncu --set full --clock-control none --import-source on -k "regex:k_trace|k_tail" -s 8 -c 18 -o gpurun_out/${TAG}_trace -f python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out | grep ${TAG}
