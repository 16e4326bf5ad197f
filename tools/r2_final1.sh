#!/bin/bash
# One-GPU evidence session of round 2: parity suite, smoke, the bench line as the driver runs it (+ a long run), the ncu launch list
# of the same command (full captures: tools/r2_ncu.sh, tools/r2_ncu_shade.sh, one ncu pass per call).  gpurun --timeout 1500 -- 'bash tools/r2_final1.sh TAG'
TAG=${1:-r2g}
mkdir -p gpurun_out
S=$(date +%s)
timeout 600 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -14 gpurun_out/${TAG}_pytest.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? $(( $(date +%s)-S ))s"; cut -c1-200 gpurun_out/${TAG}_bench.json
timeout 300 python bench.py > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; echo "bench default exit $? $(( $(date +%s)-S ))s"; cut -c1-200 gpurun_out/${TAG}_bench_default.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>/dev/null; echo "reference arm exit $? $(( $(date +%s)-S ))s"; cut -c1-200 gpurun_out/${TAG}_bench_reference.json
timeout 100 python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu list exit $? $(( $(date +%s)-S ))s"
ls -la gpurun_out | grep ${TAG}
