#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/prof_knn.py ${1:-20000} ${2:-16}"
$CMD > gpurun_out/ncu_knn_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/ncu_knn_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:knn_filter_tc_kernel -s 1 -c 1 -f -o gpurun_out/r2_knn_tc_prof3 $CMD > gpurun_out/ncu_knn.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_knn_plain.log; tail -3 gpurun_out/ncu_knn.log
