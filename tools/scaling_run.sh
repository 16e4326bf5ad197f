#!/bin/bash
# usage: gpurun --gpus G -- 'bash tools/scaling_run.sh TAG N1 N2 ...'  : bench.py at each N (one rank per GPU), lines to gpurun_out/TAG_bench_nN.json
TAG=$1; shift
mkdir -p gpurun_out
for n in "$@"; do
  if [ "$n" = "1" ]; then cmd="python bench.py --gpus 1"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n"; fi
  t0=$(date +%s)
  timeout 600 $cmd --steps 200 --warmup 3 2> gpurun_out/${TAG}_n$n.err | grep '^{' > gpurun_out/${TAG}_bench_n$n.json
  echo "N=$n rc=$? wall=$(( $(date +%s) - t0 ))s"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_n$n.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["e2e"]["value"], {k: round(v, 1) for k, v in d["phases"].items() if k.endswith("mrays_s")})
except Exception as e:
    print("no line:", e); print(open("gpurun_out/${TAG}_n$n.err").read()[-1500:])
PY
done
