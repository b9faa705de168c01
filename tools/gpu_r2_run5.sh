export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --durations=8 > gpurun_out/r2_gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_gpu_tests.log
grep -v "^$" gpurun_out/r2_gpu_tests.log | tail -40
timeout 900 python bench.py --steps 4 --warmup 3 > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err; echo "bench rc=$?"
tail -c 6000 gpurun_out/r2_bench_c2.json; tail -5 gpurun_out/r2_bench_c2.err
