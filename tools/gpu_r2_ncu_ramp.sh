#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/prof_ramp.py ${1:-3}"
$CMD > gpurun_out/ncu_ramp_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:sssp_kernel -c 2 -f -o gpurun_out/r2_sssp_ramp $CMD > gpurun_out/ncu_ramp.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/ncu_ramp_plain.log; tail -5 gpurun_out/ncu_ramp.log
