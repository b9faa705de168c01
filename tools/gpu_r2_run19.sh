mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -q -x -k "knn" 2>&1 | tail -3
for n in 256 128; do GEOCB_KNN_TC_N=$n timeout 600 python tools/prof_knn.py 60000 16 2>&1 | tail -1 | sed "s/^/N=$n /"; done
