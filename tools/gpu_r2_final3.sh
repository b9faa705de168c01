#!/bin/bash
# round-end rehearsal: what the driver runs (GPU tests, smoke, default bench, reference arm)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests_final.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r2_gpu_tests_final.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_smoke.log
timeout 900 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"
python - <<PY
import json
j=json.loads(open("gpurun_out/r2_bench_final.json").read().strip().splitlines()[-1])
b=j["build"]
print("value %.1f G/s frac %.3f e2e %.1f G/s (%.1f ms) build %.2f s (knn %.2f edge %.2f) launches %s digest %s" % (j["value"]/1e9, j["roofline"]["frac"], j["e2e"]["value"]/1e9, j["e2e"]["ms_per_step"], b["total_s"], b["knn"], b["edge_len"], j["gpu_launches"], b["digest_medoids_codes_sha256"][:8]))
print("keys:", sorted(j.keys()))
PY
