timeout 600 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -q -x -k "knn" 2>&1 | tail -2
for d in 16 32; do timeout 600 python tools/prof_knn.py 60000 $d 2>&1 | tail -1; done
