#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_parity.py tests/test_gpu_pipeline.py tests/test_gpu_c2.py -x -q -m gpu > gpurun_out/r2_knn5_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2_knn5_tests.log
{ echo "# default (one A tile per CTA, two CTAs per SM)"; python tools/prof_knn_trace.py 60000 16; echo "# GEOCB_KNN_TC_PAIR=0 (two A tiles per CTA, one CTA per SM)"; GEOCB_KNN_TC_PAIR=0 python tools/prof_knn_trace.py 60000 16; echo "# d=32 (two-tile CTA)"; python tools/prof_knn_trace.py 20000 32; } > gpurun_out/r2_knn_phase_clocks.txt 2>&1
cat gpurun_out/r2_knn_phase_clocks.txt
timeout 600 python bench.py --skip-cpu-baseline --skip-e2e > gpurun_out/r2_knn5_bench.json 2>/dev/null; python - <<PY
import json
j=json.loads(open("gpurun_out/r2_knn5_bench.json").read().strip().splitlines()[-1])
b=j["build"]; print({k:b[k] for k in ("total_s","h2d","knn","edge_len","lcc","kmedoids")}, j["value"]/1e9, b["digest_medoids_codes_sha256"][:8])
PY
