#!/bin/bash
# N-GPU session: correctness of the exchange + framebuffer gather (tools/multi_gpu_check.py), then the bench line as the driver runs it.
# gpurun --gpus N --timeout 900 -- 'bash tools/r2_multi.sh N TAG'
N=$1; TAG=${2:-r2h}
mkdir -p gpurun_out
S=$(date +%s)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 tools/multi_gpu_check.py --frames 40 > gpurun_out/${TAG}_check_n$N.json 2> gpurun_out/${TAG}_check_n$N.err; echo "check exit $? $(( $(date +%s)-S ))s"; tail -1 gpurun_out/${TAG}_check_n$N.json | cut -c1-900
timeout 400 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench exit $? $(( $(date +%s)-S ))s"; tail -1 gpurun_out/${TAG}_bench_n$N.json | cut -c1-300; tail -3 gpurun_out/${TAG}_bench_n$N.err
timeout 400 $TR --master-port 29513 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/${TAG}_bench200_n$N.json 2> gpurun_out/${TAG}_bench200_n$N.err; echo "bench200 exit $? $(( $(date +%s)-S ))s"; tail -1 gpurun_out/${TAG}_bench200_n$N.json | cut -c1-300
python - <<P
import json
for f in ("gpurun_out/${TAG}_bench_n$N.json", "gpurun_out/${TAG}_bench200_n$N.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "value %.0f e2e %.0f blocking %.0f training %.4f steady %.4f" % (d["value"], d["e2e"]["value"], d["e2e"]["blocking"]["value"], d["phases"]["training_ms_per_step"], d["phases"]["steady_ms_per_step"]), d["exchange_check"], d["e2e"]["image_complete"], d["value_windows"], d["e2e"].get("delivery"))
    except Exception as e:
        print(f, "unreadable", e)
P
