#!/bin/bash
# Build sssp.cu variants on the GPU box and time each (usage: tools/sweep_sssp.sh "PER:MB:THREADS ..." SLOTS)
cd "$(dirname "$0")/.."
python tools/prof_sssp.py 60000 4 1 > /dev/null 2>&1   # warm: import torch, build + cache the graph
for cfg in $1; do
  IFS=: read per mb th <<< "$cfg"
  rm -f vqvae_b200/csrc/build/sssp.o
  make -C vqvae_b200/csrc EXTRA="-DGEOCB_SSSP_PER=$per -DGEOCB_SSSP_MINBLOCKS=$mb -DGEOCB_SSSP_THREADS=$th" > /dev/null 2>&1
  PROF_TAG="per=$per mb=$mb th=$th" timeout 120 python tools/prof_sssp.py 60000 $2 4 2>&1 | grep -E "slots=|FAILED"
done
rm -f vqvae_b200/csrc/build/sssp.o; make -C vqvae_b200/csrc > /dev/null 2>&1
