#!/bin/bash
# One GPU-box session at the end of a round: parity tests, smoke, the bench line, the ncu launch list of the same command and
# full captures of the traversal / tail / exchange kernels.  Run as: gpurun --timeout 900 -- 'bash tools/gpu_final.sh TAG'
TAG=${1:-r3f}
mkdir -p gpurun_out
S=$(date +%s)
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 200 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? $(( $(date +%s)-S ))s"; cut -c1-300 gpurun_out/${TAG}_bench.json
timeout 100 python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu list exit $? $(( $(date +%s)-S ))s"
timeout 200 ncu --set full --clock-control none --import-source on -k "regex:k_trace|k_tail" -s 17 -c 5 -o gpurun_out/${TAG}_trace -f python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full exit $? $(( $(date +%s)-S ))s"
timeout 60 python tools/peer_profile.py 6 > gpurun_out/${TAG}_peer_plain.log 2>&1 &&
timeout 200 ncu --set full --clock-control none --import-source on -k "regex:k_peer|k_tail|k_rc_update" -s 8 -c 10 -o gpurun_out/${TAG}_peer -f python tools/peer_profile.py 6 > gpurun_out/${TAG}_ncu3.log 2>&1
echo "ncu peer exit $? $(( $(date +%s)-S ))s"
ls -la gpurun_out | tail -20
