#!/bin/bash
# One GPU-box session: parity tests, smoke, a short bench line, the ncu launch list and a full capture of
# the traversal kernels.  Run as: gpurun --timeout 1500 -- 'bash tools/gpu_check.sh [tag]'
TAG=${1:-r1}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${TAG}_smoke.log
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json
python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu list exit $?"
ncu --set full --clock-control none --import-source on -k "regex:k_trace|k_tail" -s 17 -c 5 -o gpurun_out/${TAG}_trace -f python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
ls -la gpurun_out
