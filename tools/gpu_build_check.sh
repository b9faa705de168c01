#!/bin/bash
# edge-length parity tests, then the build-stage seconds at 1 GPU (and N GPUs when $1 is given)
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null
timeout 300 python -m pytest tests -x -q -m gpu -k "edge or riemann or pipeline" > gpurun_out/r1b_tests.log 2>&1 || { echo "tests failed"; tail -15 gpurun_out/r1b_tests.log; exit 1; }
tail -1 gpurun_out/r1b_tests.log
timeout 200 python bench.py --steps 3 --warmup 3 --skip-e2e --skip-cpu-baseline > gpurun_out/r1b_1gpu.json 2> gpurun_out/r1b_1gpu.err || { echo "bench failed"; tail -5 gpurun_out/r1b_1gpu.err; exit 1; }
python -c "import json; d=json.loads([l for l in open('gpurun_out/r1b_1gpu.json') if l.startswith('{')][-1]); print(1, d['value']/1e9, d['build'])"
if [ -n "$1" ]; then
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $1 --steps 3 --warmup 3 --skip-e2e > gpurun_out/r1b_ngpu.json 2> gpurun_out/r1b_ngpu.err || { echo "bench N failed"; tail -5 gpurun_out/r1b_ngpu.err; exit 1; }
python -c "import json; d=json.loads([l for l in open('gpurun_out/r1b_ngpu.json') if l.startswith('{')][-1]); print($1, d['value']/1e9, d['build'])"
head -c 300 gpurun_out/r1b_ngpu.json
fi
