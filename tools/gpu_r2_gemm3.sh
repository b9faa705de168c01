for m in 0 1 2 3; do
  echo "== GEOCB_GEMM3X_MODE=$m"
  GEOCB_GEMM3X_MODE=$m timeout 600 python -m pytest tests/test_gpu_jvp_gemm.py -q -s 2>&1 | grep -E "^.?M=.*gemm3x err" | cut -c1-100
  GEOCB_GEMM3X_MODE=$m timeout 600 python tools/edge_len_noise.py 2>&1 | grep "^tc" | cut -c1-200
done
