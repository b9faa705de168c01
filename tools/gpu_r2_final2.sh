#!/bin/bash
# full GPU suite + the contract bench (with CPU baselines) + diag of the lower-level e2e path
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests_final.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r2_gpu_tests_final.log
timeout 300 python tools/diag_e2e_dev.py > gpurun_out/r2_diag_e2e_dev.log 2>&1; echo "diag rc=$?"; cat gpurun_out/r2_diag_e2e_dev.log
timeout 1500 python bench.py > gpurun_out/r2_bench_c2_b.json 2> gpurun_out/r2_bench_c2_b.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r2_bench_c2_b.json
