#!/bin/bash
# usage: gpurun -- 'bash tools/shard_sweep.sh tag WORLD "A=1" "B=2" ...'   ("-" = no setting): one rank's share per env setting
TAG=$1; WORLD=$2; shift; shift
mkdir -p gpurun_out
for e in "$@"; do
  echo "== world $WORLD env $e" | tee -a gpurun_out/${TAG}_shard.log
  [ "$e" = "-" ] && e=""
  env $e python tools/shard_probe.py $WORLD 100 2>&1 | tail -5 | tee -a gpurun_out/${TAG}_shard.log
done
