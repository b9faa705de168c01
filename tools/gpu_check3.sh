#!/bin/bash
# cosine kNN parity test, then a 2-GPU bench whose stdout must be exactly one JSON line
mkdir -p gpurun_out
python -c "import torch" 2>/dev/null
timeout 200 python -m pytest tests -x -q -m gpu -k "knn" > gpurun_out/r1c_tests.log 2>&1 || { echo "tests failed"; tail -25 gpurun_out/r1c_tests.log; }
tail -1 gpurun_out/r1c_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r1c_2gpu.json 2> gpurun_out/r1c_2gpu.err || { echo "bench failed"; tail -5 gpurun_out/r1c_2gpu.err; exit 1; }
echo "stdout lines: $(wc -l < gpurun_out/r1c_2gpu.json)"; head -c 200 gpurun_out/r1c_2gpu.json; echo
python -c "import json; d=json.loads(open('gpurun_out/r1c_2gpu.json').read()); print(d['value']/1e9, d['e2e'], d['build'])"
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/r1c_ref.json 2> gpurun_out/r1c_ref.err; echo "ref rc=$? lines $(wc -l < gpurun_out/r1c_ref.json)"; head -c 600 gpurun_out/r1c_ref.json
