export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py -q -x -k "knn or medoid" > gpurun_out/r2_knn_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_knn_tests.log
grep -v "^$" gpurun_out/r2_knn_tests.log | tail -40
timeout 900 python bench.py --steps 4 --warmup 3 > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err; echo "bench rc=$?"
tail -c 7000 gpurun_out/r2_bench_c2.json; grep -v "^frame" gpurun_out/r2_bench_c2.err | tail -8
