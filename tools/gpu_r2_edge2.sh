#!/bin/bash
mkdir -p gpurun_out
for mode in 0 blas 0 blas; do GEOCB_OVERLAP_DECODER_SETUP=$mode timeout 600 python bench.py --skip-cpu-baseline --skip-e2e > gpurun_out/r2_edge2_bench_$mode.json 2>/dev/null; python - <<PY
import json
j=json.loads(open("gpurun_out/r2_edge2_bench_$mode.json").read().strip().splitlines()[-1])
b=j["build"]; print("overlap=$mode", {k:b[k] for k in ("total_s","h2d","knn","edge_len","lcc","kmedoids")}, b["digest_medoids_codes_sha256"][:8])
PY
done
