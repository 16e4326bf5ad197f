#!/bin/bash
# N-GPU bench line only (the driver's command), when GPU minutes are short.  gpurun --gpus N -- 'bash tools/r2_multi_slim.sh N TAG'
N=$1; TAG=${2:-r2w}
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench exit $?"
python - <<P
import json
d = json.loads([l for l in open("gpurun_out/${TAG}_bench_n$N.json") if l.startswith("{")][-1])
print("value %.0f e2e %.0f training %.4f steady %.4f" % (d["value"], d["e2e"]["value"], d["phases"]["training_ms_per_step"], d["phases"]["steady_ms_per_step"]), d["exchange_check"]["replicated"], d["e2e"]["image_complete"], {k: (round(v["value"]) if isinstance(v, dict) and v.get("value") else v) for k, v in d["e2e"]["delivery"].items()})
P
