export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err; echo "c2 rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/r2_bench_c2_reference.json 2> gpurun_out/r2_bench_c2_reference.err; echo "ref rc=$?"; tail -c 600 gpurun_out/r2_bench_c2_reference.json
bash tools/gpu_r2_workloads.sh c3 c5:1e6,64,4096 c5:4e6,16,512 c5:1.6e7,16,512
