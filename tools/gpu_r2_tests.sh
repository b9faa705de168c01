export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --durations=15 > gpurun_out/r2_gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_gpu_tests.log
tail -60 gpurun_out/r2_gpu_tests.log
