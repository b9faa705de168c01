mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_jvp_gemm.py -q -s 2>&1 | grep -E "gemm3x err|passed|failed"
timeout 600 python tools/edge_len_noise.py 2>&1 | tail -4
