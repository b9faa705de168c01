#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_jvp_gemm.py tests/test_gpu_pipeline.py tests/test_gpu_legacy.py tests/test_gpu_ref_conformance.py -x -q -m gpu > gpurun_out/r2_edge1_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2_edge1_tests.log
python tools/prof_edge_len.py 2>&1 | grep -v Warn | cut -c1-72,96-110,166-240 | head -16
GEOCB_JVP_FUSE_IN=0 python tools/prof_edge_len.py 2>&1 | head -1
