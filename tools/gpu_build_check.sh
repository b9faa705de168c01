#!/bin/bash
# edge-length parity tests, then the build-stage seconds at 1 GPU (and N GPUs when $1 is given)
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null




if [ -n "$1" ]; then
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $1 --steps 3 --warmup 3 --skip-e2e > gpurun_out/r1b_ngpu.json 2> gpurun_out/r1b_ngpu.err || { echo "bench N failed"; tail -5 gpurun_out/r1b_ngpu.err; exit 1; }
python -c "import json; d=json.loads([l for l in open('gpurun_out/r1b_ngpu.json') if l.startswith('{')][-1]); print($1, d['value']/1e9, d['build'])"
head -c 300 gpurun_out/r1b_ngpu.json
fi
