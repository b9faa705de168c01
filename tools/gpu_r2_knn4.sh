#!/bin/bash
# kNN filter, two one-tile CTAs per SM (GEOCB_KNN_TC_PAIR=1) against the default two-tile CTA
mkdir -p gpurun_out
GEOCB_KNN_TC_PAIR=1 timeout 900 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -x -q -m gpu -k "knn" > gpurun_out/r2_knn4_tests.log 2>&1; echo "tests(pair) rc=$?"; tail -3 gpurun_out/r2_knn4_tests.log
{ echo "# default (Cfg16: 2 A tiles per CTA, 1 CTA/SM)"; python tools/prof_knn.py 60000 16; echo "# GEOCB_KNN_TC_PAIR=1 (Cfg16P: 1 A tile per CTA, 2 CTAs/SM)"; GEOCB_KNN_TC_PAIR=1 python tools/prof_knn.py 60000 16; GEOCB_KNN_TC_PAIR=1 python tools/prof_knn_trace.py 60000 16; } > gpurun_out/r2_knn4_prof.log 2>&1
cat gpurun_out/r2_knn4_prof.log
