#!/bin/bash
# kNN filter with FMNMX3 min-trees + octet survivor search: parity tests, then timing at C2 size and d = 32 / 64
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -x -q -m gpu -k "knn" > gpurun_out/r2_knn3_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2_knn3_tests.log
timeout 600 python tools/prof_knn.py 60000 16 > gpurun_out/r2_knn3_prof.log 2>&1; timeout 300 python tools/prof_knn.py 20000 32 >> gpurun_out/r2_knn3_prof.log 2>&1; timeout 300 python tools/prof_knn.py 20000 64 >> gpurun_out/r2_knn3_prof.log 2>&1
cat gpurun_out/r2_knn3_prof.log
