#!/bin/bash
# ncu --set full of the first sssp_kernel launch of bench.py's timed region (one streamed 64-source launch)
mkdir -p gpurun_out
CMD="python bench.py --skip-cpu-baseline --skip-e2e"
$CMD > gpurun_out/ncufull_plain.json 2> gpurun_out/ncufull_plain.err || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:sssp_kernel -c 1 -f -o gpurun_out/r2_sssp_bench_launch $CMD > gpurun_out/ncufull.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncufull.log | cut -c1-300
