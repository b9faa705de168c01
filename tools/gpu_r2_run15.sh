mkdir -p gpurun_out
for ov in 1 0 1 0; do
GEOCB_OVERLAP_DECODER_SETUP=$ov timeout 900 python bench.py --steps 1 --warmup 3 --skip-cpu-baseline --skip-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); b=d['build']
print('overlap=$ov', {k:round(v,3) for k,v in b.items() if k in('total_s','h2d','knn','edge_len','kmedoids')})"
done
./tools/ubench/tmem_ld > gpurun_out/r2_ubench_tmem_ld.txt 2>&1; cat gpurun_out/r2_ubench_tmem_ld.txt
