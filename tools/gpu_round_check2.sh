#!/bin/bash
# SSSP parity tests, a plain short bench, then one `ncu --set full` capture of the first sssp_kernel launch inside
# the bench's timed region (DRAM traffic for roofline.traffic).  Stops at the first failure.
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null
timeout 300 python -m pytest tests -x -q -m gpu -k "sssp or kmedoids or pipeline" > gpurun_out/r1y_tests.log 2>&1 || { echo "tests failed"; tail -15 gpurun_out/r1y_tests.log; exit 1; }
tail -1 gpurun_out/r1y_tests.log
timeout 240 python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-e2e > gpurun_out/r1y_plain.json 2> gpurun_out/r1y_plain.err || exit 1
timeout 400 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:sssp_kernel -c 1 \
  -o gpurun_out/r1y_sssp_bench -f python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-e2e > gpurun_out/r1y_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r1y_ncu.log
