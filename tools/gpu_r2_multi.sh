# usage: bash tools/gpu_r2_multi.sh N
N=${1:-2}
export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/r2_multi_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_multi_tests.log
  grep -v "^$" gpurun_out/r2_multi_tests.log | tail -30
fi
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 4 --warmup 3 > gpurun_out/r2_bench_c2_${N}gpu.json 2> gpurun_out/r2_bench_c2_${N}gpu.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r2_bench_c2_${N}gpu.json; grep -v "^frame" gpurun_out/r2_bench_c2_${N}gpu.err | tail -12
