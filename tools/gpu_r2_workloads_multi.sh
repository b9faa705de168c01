#!/bin/bash
# usage: bash tools/gpu_r2_workloads_multi.sh N workload...   (bench.py --workload under torchrun on N GPUs)
N=$1; shift
export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
port=29520
for w in "$@"; do
  tag=$(echo "$w" | tr ':,' '__')
  port=$((port + 1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 3 --warmup 3 --workload "$w" > gpurun_out/r2_bench_${tag}_${N}gpu.json 2> gpurun_out/r2_bench_${tag}_${N}gpu.err
  echo "$w rc=$?"
  python - <<PY
import json
try:
    j = json.loads(open("gpurun_out/r2_bench_${tag}_${N}gpu.json").read().strip().splitlines()[-1])
    b = j["build"]
    print(j["config"]["workload"], "| n", j["config"]["n"], "| value %.1f G/s frac %.3f" % (j["value"] / 1e9, j["roofline"]["frac"]),
          "| build", {k: round(b[k], 3) for k in ("total_s", "knn", "edge_len", "lcc", "kmedoids") if k in b}, "| e2e %.1f G/s" % (j["e2e"]["value"] / 1e9),
          "| assign_sharded", (j.get("assign_sharded") or {}).get("ms"))
except Exception as e:
    print("no line:", e)
PY
done
