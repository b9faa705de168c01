#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests_final.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2_gpu_tests_final.log
{ echo "# default: bf16-split operands, one A tile per CTA, two CTAs per SM"; python tools/prof_knn_trace.py 60000 16; echo "# GEOCB_KNN_TC_FMT=tf32 (one A tile per CTA, two CTAs per SM)"; GEOCB_KNN_TC_FMT=tf32 python tools/prof_knn_trace.py 60000 16; echo "# GEOCB_KNN_TC_FMT=tf32 GEOCB_KNN_TC_PAIR=0 (two A tiles per CTA, one CTA per SM: the configuration up to this round)"; GEOCB_KNN_TC_FMT=tf32 GEOCB_KNN_TC_PAIR=0 python tools/prof_knn_trace.py 60000 16; echo "# d=32, bf16 / tf32"; python tools/prof_knn_trace.py 20000 32; GEOCB_KNN_TC_FMT=tf32 python tools/prof_knn_trace.py 20000 32; } > gpurun_out/r2_knn_phase_clocks.txt 2>&1
cat gpurun_out/r2_knn_phase_clocks.txt | grep -v "^  "
timeout 600 python bench.py --skip-cpu-baseline --skip-e2e > gpurun_out/r2_knn7_bench.json 2>/dev/null; timeout 600 python bench.py --workload c3 --skip-cpu-baseline --skip-e2e > gpurun_out/r2_knn7_bench_c3.json 2>/dev/null; python - <<PY
import json
for f in ("r2_knn7_bench","r2_knn7_bench_c3"):
    j=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    b=j["build"]; print(f, {k:b[k] for k in ("total_s","h2d","knn","edge_len","lcc","kmedoids")}, j["value"]/1e9, b["digest_medoids_codes_sha256"][:8])
PY
