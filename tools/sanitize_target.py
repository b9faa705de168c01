"""Small end-to-end workload for compute-sanitizer: smoke() (kNN -> SSSP -> K-medoids on 2000 latents) plus the s1
pipeline (kNN, tcgen05 filter, edge lengths incl. the tcgen05 GEMM and the fused BatchNorm-JVP kernel, LCC, K-medoids)
and one vertex-major SSSP call.  Usage: compute-sanitizer --tool memcheck|racecheck python tools/sanitize_target.py"""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "tests", ROOT / "oracle"):
    sys.path.insert(0, str(p))
import __graft_entry__ as entry
from conftest import s1_latents, fashion_decoder
from vqvae_b200.pipeline import build_codebook_device
from vqvae_b200 import device_graph as dg

entry.smoke()
z4, _ = s1_latents()
res = build_codebook_device(z4, fashion_decoder().cuda(), k=10, sym="union", K=32, verbose=False)
s1 = dict(np.load(ROOT / "tests" / "golden" / "s1_stages.npz"))
assert np.array_equal(res.medoids, s1["medoids"]) and np.array_equal(res.codes, s1["codes"])
D = dg.sssp_multi_vm(res.g_lcc, [0, 7, 1999, 1234], blocks=3, blk_min_active=0)
assert torch.equal(D, dg.sssp_multi(res.g_lcc, [0, 7, 1999, 1234], kernel="slot"))
torch.cuda.synchronize()
print("sanitize target ok", flush=True)
