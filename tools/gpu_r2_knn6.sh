#!/bin/bash
# bf16 operand format of the kNN filter (GEOCB_KNN_TC_FMT=bf16) against tf32
mkdir -p gpurun_out
GEOCB_KNN_TC_FMT=bf16 timeout 900 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -x -q -m gpu -k "knn or medoid" > gpurun_out/r2_knn6_tests.log 2>&1; echo "tests(bf16) rc=$?"; tail -3 gpurun_out/r2_knn6_tests.log
{ for fmt in tf32 bf16; do echo "# GEOCB_KNN_TC_FMT=$fmt"; GEOCB_KNN_TC_FMT=$fmt python tools/prof_knn.py 60000 16; GEOCB_KNN_TC_FMT=$fmt python tools/prof_knn.py 20000 32; GEOCB_KNN_TC_FMT=$fmt python tools/prof_knn.py 20000 64; done; GEOCB_KNN_TC_FMT=bf16 python tools/prof_knn_trace.py 60000 16; } > gpurun_out/r2_knn6_prof.log 2>&1
cat gpurun_out/r2_knn6_prof.log
