#!/bin/bash
# staged check of the streamed-source SSSP: parity first, then throughput at several slot / source counts
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null
for st in 0 1 2; do
GEOCB_SSSP_STAGGER=$st timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sssp" > gpurun_out/stream_tests_$st.log 2>&1
echo "stagger=$st tests rc=$?" | tee -a gpurun_out/stream_tests_$st.log
tail -3 gpurun_out/stream_tests_$st.log
grep -q "rc=0" gpurun_out/stream_tests_$st.log || exit 1
done
: > gpurun_out/stream_prof.log
for cfg in "0 16 64" "1 16 64" "2 16 64" "1 24 64" "0 16 128" "1 16 128" "1 24 128" "1 32 128" "0 16 512" "1 16 512" "2 16 512" "1 24 512" "1 32 512"; do
  set -- $cfg
  GEOCB_SSSP_STAGGER=$1 PROF_TAG=stagger$1 timeout 600 python tools/prof_sssp.py 60000 $2 3 --sources $3 >> gpurun_out/stream_prof.log 2>&1
done
cat gpurun_out/stream_prof.log
