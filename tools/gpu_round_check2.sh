#!/bin/bash
# round check 2: GPU parity suite, the default bench line, then one `ncu --set full` capture of the first sssp_kernel
# launch inside the bench's timed region (DRAM traffic for roofline.traffic).  Stops at the first failure.
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/r1y_tests.log 2>&1 || { echo "tests failed"; tail -15 gpurun_out/r1y_tests.log; exit 1; }
tail -2 gpurun_out/r1y_tests.log
timeout 420 python bench.py > gpurun_out/r1y_bench.json 2> gpurun_out/r1y_bench.err || { echo "bench failed"; tail -5 gpurun_out/r1y_bench.err; exit 1; }
cat gpurun_out/r1y_bench.json
timeout 240 python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-e2e > gpurun_out/r1y_plain.json 2> gpurun_out/r1y_plain.err || exit 1
timeout 500 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:sssp_kernel -c 1 \
  -o gpurun_out/r1y_sssp_bench -f python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-e2e > gpurun_out/r1y_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r1y_ncu.log
