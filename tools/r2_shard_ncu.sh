#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/shard_launches.py 8 6 > gpurun_out/shard8_now.log 2>&1 || exit 1
tail -n 2 gpurun_out/shard8_now.log | cut -c1-600
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/shard8_launches_now.csv python tools/shard_launches.py 8 6 > gpurun_out/shard8_ncu.log 2>&1
echo ncu rc=$?
