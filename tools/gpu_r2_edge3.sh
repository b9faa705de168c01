#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_jvp_gemm.py tests/test_gpu_pipeline.py tests/test_gpu_legacy.py tests/test_gpu_ref_conformance.py -x -q -m gpu > gpurun_out/r2_edge3_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2_edge3_tests.log
python tools/prof_edge_len.py 2>&1 | grep -v Warn | cut -c1-72,96-110,166-240 | head -12 | tee gpurun_out/r2_edge3_prof.log
for i in 1 2; do timeout 600 python bench.py --skip-cpu-baseline --skip-e2e > gpurun_out/r2_edge3_bench_$i.json 2>/dev/null; python - <<PY
import json
j=json.loads(open("gpurun_out/r2_edge3_bench_$i.json").read().strip().splitlines()[-1])
b=j["build"]; print({k:b[k] for k in ("total_s","h2d","knn","edge_len","lcc","kmedoids")}, b["digest_medoids_codes_sha256"][:8])
PY
done
