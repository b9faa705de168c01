timeout 300 python tools/prof_edge_fixed.py 2>&1 | tail -9
for lib in vqvae_b200/libgeocb.so vqvae_b200/libgeocb_knncoop.so; do
  echo "== $lib"
  GEOCB_LIB=$PWD/$lib timeout 600 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -q -x -k "knn" 2>&1 | tail -1
  for d in 16 32; do GEOCB_LIB=$PWD/$lib timeout 600 python tools/prof_knn.py 60000 $d 2>&1 | tail -1; done
done
