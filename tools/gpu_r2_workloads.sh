# usage: bash tools/gpu_r2_workloads.sh "c3" "c5:1e6,64,4096" ...   (1 GPU each, short step counts)
mkdir -p gpurun_out
for w in "$@"; do
  tag=$(echo "$w" | tr ':,' '___')
  timeout 1500 python bench.py --workload "$w" --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/r2_bench_${tag}.json 2> gpurun_out/r2_bench_${tag}.err; echo "$w rc=$?"
  python - "$tag" <<'PY'
import json,sys
try:
    d=json.loads(open(f"gpurun_out/r2_bench_{sys.argv[1]}.json").read().strip().splitlines()[-1])
    b=d["build"]
    print(d["config"]["workload"], "| n",d["config"]["n"],"nnz",d["config"]["nnz"],"| value %.1f G/s frac %.3f"%(d["value"]/1e9,d["roofline"]["frac"]),
          "| build", {k:(round(v,3) if isinstance(v,float) else v) for k,v in b.items() if k in("total_s","knn","edge_len","kmedoids","lcc")},
          "| e2e %.1f G/s"%(d["e2e"]["value"]/1e9))
except Exception as e:
    print("parse failed", e)
PY
  grep -v "^frame" gpurun_out/r2_bench_${tag}.err | tail -4
done
