#!/bin/bash
# vm32 kernel: parity tests, then the C2-graph sweep against the slot-major kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sssp_vm.py -x -q -m gpu > gpurun_out/r2_vm32_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_vm32_tests.log
timeout 1200 python tools/sweep_vm.py --gen 32,16 --blocks ${BLOCKS:-1,8} --scales ${SCALES:-4} --minact 16 --trace > gpurun_out/r2_sweep_vm32.log 2>&1
echo "sweep rc=$?"; cat gpurun_out/r2_sweep_vm32.log | cut -c1-900
