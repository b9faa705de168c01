timeout 600 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -q -x -k "knn" 2>&1 | tail -1
GEOCB_KNN_TC_ES=2 timeout 600 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -q -x -k "knn" 2>&1 | tail -1
for rep in 1 2; do
GEOCB_LIB=$PWD/vqvae_b200/libgeocb_prev.so timeout 600 python tools/prof_knn.py 60000 16 2>&1 | tail -1 | sed "s/^/prev /"
for es in 1 2; do GEOCB_KNN_TC_ES=$es timeout 600 python tools/prof_knn.py 60000 16 2>&1 | tail -1 | sed "s/^/new ES=$es /"; done
done
timeout 600 python tools/prof_knn.py 60000 32 2>&1 | tail -1
