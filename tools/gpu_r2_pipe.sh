#!/bin/bash
# streamed SSSP launch: parity tests of everything that touches sssp.cu, then the slots x stagger sweep on the C2 graph
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sssp_vm.py tests/test_gpu_c2.py -x -q -m gpu > gpurun_out/r2_pipe_sssp_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_pipe_sssp_tests.log
timeout 1200 python tools/sweep_pipe.py --slots ${SLOTS:-16,24,32} --stagger ${STAGGER:-0,8,12,16} --trace > gpurun_out/r2_sweep_pipe.log 2>&1
echo "sweep rc=$?"; cut -c1-1500 gpurun_out/r2_sweep_pipe.log
