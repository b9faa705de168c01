#!/bin/bash
# launch list of bench.py's timed region (cudaProfilerStart/Stop bracket it), after the same command exited 0 without ncu
mkdir -p gpurun_out
CMD="python bench.py --skip-cpu-baseline --skip-e2e"
$CMD > gpurun_out/launchlist_plain.json 2> gpurun_out/launchlist_plain.err || { echo "plain run failed"; tail -5 gpurun_out/launchlist_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file gpurun_out/r2_bench_launches.csv $CMD > gpurun_out/launchlist_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/launchlist_ncu.log | cut -c1-300; grep -c sssp_kernel gpurun_out/r2_bench_launches.csv; tail -8 gpurun_out/r2_bench_launches.csv | cut -c1-250
