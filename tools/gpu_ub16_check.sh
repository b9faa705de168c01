#!/bin/bash
# default build and the GEOCB_SSSP_UB16 variant: SSSP / K-medoids / pipeline parity, then throughput on the C2 graph
make -C vqvae_b200/csrc ub16 > /dev/null 2>&1
V=$PWD/vqvae_b200/libgeocb_ub16.so
python -c "import torch" 2>/dev/null
timeout 200 python -m pytest tests -x -q -m gpu -k "sssp or kmedoids or pipeline" 2>&1 | tail -1
GEOCB_LIB=$V timeout 200 python -m pytest tests -x -q -m gpu -k "sssp or kmedoids or pipeline" 2>&1 | tail -3
timeout 240 python tools/sweep_delta.py 4 8,16,24,32 $V 2>&1 | tail -10
