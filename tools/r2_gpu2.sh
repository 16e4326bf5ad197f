#!/bin/bash
# Second GPU session: full parity suite, cooperative traversal (variant library) parity + A/B, queue staging A/B.
TAG=${1:-r2b}
mkdir -p gpurun_out
S=$(date +%s)
timeout 900 python -m pytest tests -m gpu -q --durations=12 > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? $(( $(date +%s)-S ))s"; tail -40 gpurun_out/${TAG}_pytest.log
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_coop.so
echo "== coop parity (everything cooperative: PTB200_COOP=100000)"
PTB200_COOP=100000 timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_edge_cases.py tests/test_gpu_vs_reference_device.py "tests/test_gpu_baseline_configs.py::test_c1_teapot_640x480_known_answers_on_the_gpu" -q -x > gpurun_out/${TAG}_pytest_coop.log 2>&1; echo "coop pytest exit $? $(( $(date +%s)-S ))s"; tail -15 gpurun_out/${TAG}_pytest_coop.log
echo "== coop parity (PTB200_COOP=2)"
PTB200_COOP=2 timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_vs_reference_device.py -q -x > gpurun_out/${TAG}_pytest_coop2.log 2>&1; echo "coop2 pytest exit $? $(( $(date +%s)-S ))s"; tail -5 gpurun_out/${TAG}_pytest_coop2.log
timeout 400 python tools/ab_probe.py base= coop1=PTB200_COOP=1 coop2=PTB200_COOP=2 coop4=PTB200_COOP=4 coop8=PTB200_COOP=8 coop2wave=PTB200_COOP=2,PTB200_TAIL_BOUNCE=8 coop4wave=PTB200_COOP=4,PTB200_TAIL_BOUNCE=8 coop2nopair=PTB200_COOP=2,PTB200_NO_TAIL_PAIR=1 pdl_always=PTB200_PDL_ALWAYS=1 > gpurun_out/${TAG}_ab_coop.log 2>&1; echo "ab exit $? $(( $(date +%s)-S ))s"; cat gpurun_out/${TAG}_ab_coop.log
export PTB200_LIB=$PWD/cuda_pathtracer_b200/libptb200_bulk.so
timeout 200 python tools/ab_probe.py bulk= > gpurun_out/${TAG}_ab_bulk.log 2>&1; echo "bulk ab exit $? $(( $(date +%s)-S ))s"; cat gpurun_out/${TAG}_ab_bulk.log
timeout 200 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -q -x 2>&1 | tail -3
