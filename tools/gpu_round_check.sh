#!/bin/bash
# round check: GPU parity suite, the default bench line, then the ncu launch list of the bench's timed region.
# Every stage has its own short timeout and the script stops at the first failure.
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null
timeout 240 python -m pytest tests -x -q -m gpu -k "sssp" > gpurun_out/r1x_tests_sssp.log 2>&1 || { echo "sssp tests failed"; tail -5 gpurun_out/r1x_tests_sssp.log; exit 1; }
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/r1x_tests.log 2>&1 || { echo "tests failed"; tail -15 gpurun_out/r1x_tests.log; exit 1; }
tail -2 gpurun_out/r1x_tests.log
timeout 420 python bench.py > gpurun_out/r1x_bench.json 2> gpurun_out/r1x_bench.err || { echo "bench failed"; tail -5 gpurun_out/r1x_bench.err; exit 1; }
cat gpurun_out/r1x_bench.json
timeout 240 python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-e2e > gpurun_out/r1x_plain.json 2> gpurun_out/r1x_plain.err || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/r1x_launches.csv python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-e2e > gpurun_out/r1x_ncu.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/r1x_launches.csv
