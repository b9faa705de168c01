"""Multi-GPU parity: torchrun with 2 ranks over NCCL -- the query-/batch-sharded build and the
source-sharded K-medoids calls must reproduce the single-GPU medoids, codes, QE and edge weights bit for bit.
Skipped on a box with fewer than two GPUs (the gloo world-2 tests cover the host logic on CPU)."""
import json
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
def test_two_rank_build_and_sharded_assignment_match_single_gpu():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29417", str(ROOT / "tests" / "multi_gpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, cwd=str(ROOT), timeout=900)
    assert p.returncode == 0, p.stderr[-3000:]
    line = [ln for ln in p.stdout.splitlines() if ln.startswith("MULTI_GPU_RESULT ")][-1]
    res = json.loads(line[len("MULTI_GPU_RESULT "):])
    assert res["world"] == 2 and res["all_ranks_ok"], res
    assert res["build_equal"] and res["random_init_equal"] and res["assign_equal"] and res["qe_equal"], res
    assert res["sharded_upload_equal"] and res["sssp_on_sharded_equal"], res
