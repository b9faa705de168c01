"""ncu target: the slot-major SSSP kernel on the C2 geodesic graph, 16 sources in 16 tables, three launches:
  1. warm-up, all rounds;  2. rounds 0..6 only (the narrow ramp-up rounds; GEOCB_SSSP_MAX_ROUNDS cuts the launch
  short, the call reports non-convergence, which is expected here);  3. all rounds.
Profile with `ncu --profile-from-start off -k regex:sssp_kernel -c 2`: launch 2 is the ramp phase alone, launch 3 minus launch 2 the
wide rounds and the tail."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from vqvae_b200 import device_graph as dg  # noqa: E402
from vqvae_b200.models import SpatialVAE  # noqa: E402
from vqvae_b200.pipeline import build_codebook_device  # noqa: E402

dev = torch.device("cuda")
z4 = torch.randn(60000, 16, 7, 7, generator=torch.Generator().manual_seed(42))
torch.manual_seed(0)
decoder = SpatialVAE(**bench.FASHION_CFG).decoder.to(dev)
res = build_codebook_device(z4, decoder, k=20, sym="union", K=512, init="kpp", seed=42, batch_size=512, verbose=False)
g = res.g_lcc
src = np.asarray(res.medoids, dtype=np.int64)[:16]
D = torch.empty((16, g.n), dtype=torch.float32, device=dev)
dg.sssp_multi(g, src, slots=16, out=D, kernel="slot", stagger=0)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()  # (the codebook build above launches sssp_kernel ~1500 times: Voronoi inserts)
os.environ["GEOCB_SSSP_MAX_ROUNDS"] = sys.argv[1] if len(sys.argv) > 1 else "3"  # x2 waves = rounds 0..6
try:
    dg.sssp_multi(g, src, slots=16, out=D, kernel="slot", stagger=0)
except Exception as e:
    print("cut short as intended:", str(e)[:100])
del os.environ["GEOCB_SSSP_MAX_ROUNDS"]
_, st = dg.sssp_multi(g, src, slots=16, out=D, kernel="slot", stagger=0, stats=True)
torch.cuda.synchronize()
print("full launch:", st)
