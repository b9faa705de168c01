#!/bin/bash
# bench line (device-resident part only) under a few runtime knobs / library variants
python -c "import torch" 2>/dev/null
run() { echo "== $*"; env "$@" timeout 200 python bench.py --skip-cpu-baseline --skip-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['sssp_stats'], d['build']['total_s'])"; }
for v in "$@"; do run $v; done
