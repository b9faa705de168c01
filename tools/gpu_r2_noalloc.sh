#!/bin/bash
# the allocation-free entry point's test: alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r2_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests_final.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2_gpu_tests_final.log
