export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_jvp_gemm.py tests/test_gpu_legacy.py -q > gpurun_out/r2_pipe_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pipe_tests.log
grep -v "^$" gpurun_out/r2_pipe_tests.log | tail -12 | cut -c1-300
timeout 900 python bench.py --steps 4 --warmup 3 --skip-cpu-baseline > gpurun_out/r2_bench_c2_b.json 2> gpurun_out/r2_bench_c2_b.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2_bench_c2_b.json").read().strip().splitlines()[-1])
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in d["build"].items() if k!="stage_rooflines"})
print("value %.1f e2e %.1f"%(d["value"]/1e9, d["e2e"]["value"]/1e9))
PY
grep -v "^frame" gpurun_out/r2_bench_c2_b.err | tail -5
