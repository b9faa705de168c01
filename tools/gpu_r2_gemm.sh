mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_jvp_gemm.py -x -q -s > gpurun_out/r2_gemm_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_gemm_tests.log
grep -v "^$" gpurun_out/r2_gemm_tests.log | tail -25
timeout 600 python -m pytest tests/test_gpu_pipeline.py -x -q > gpurun_out/r2_pipe_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pipe_tests.log
grep -v "^$" gpurun_out/r2_pipe_tests.log | tail -25
timeout 600 python tools/prof_edge_len.py > gpurun_out/r2_prof_edge_len.log 2>&1; tail -30 gpurun_out/r2_prof_edge_len.log
