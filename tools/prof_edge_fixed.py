"""Where does the per-process fixed cost of the edge-length stage go?  Times the first and second call on one
32 768-edge chunk and the pieces of the first (plan, weight image, workspace, first GEMMs)."""
import sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench
from vqvae_b200.models import SpatialVAE
from vqvae_b200.geo import riemannian_metric as rm
from vqvae_b200 import device_graph as dg, _lib
_lib.load()
dev = torch.device("cuda")
torch.manual_seed(0)
decoder = SpatialVAE(**bench.FASHION_CFG).decoder.to(dev)
g = torch.Generator().manual_seed(1)
z = torch.randn(500000, 16, generator=g).to(dev)
E = 64 * 512
ei = torch.randint(0, 500000, (E,), generator=g, dtype=torch.int32).to(dev)
ej = torch.randint(0, 500000, (E,), generator=g, dtype=torch.int32).to(dev)
torch.cuda.synchronize()
def t(label, fn):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize()
    print(f"{label}: {(time.perf_counter()-t0)*1e3:.1f} ms", flush=True); return r
t("decoder_plan (unroll to dense operators)", lambda: rm.decoder_plan(decoder, 16))
t("torch.zeros 537 MB workspace", lambda: dg.gemm3x_image_buffer(131072, 512, dev))
ops = rm.decoder_plan(decoder, 16)[0]
w = [o[1] for o in ops if o[0] == "lin"]
t("first cuBLAS GEMM (handle creation)", lambda: torch.zeros(1024, 16, device=dev) @ w[0])
t("weight image", lambda: dg._gemm3x_weight_image(w[1]))
t("first edge_lengths_device call (32768 edges)", lambda: rm.edge_lengths_device(decoder, z, ei, ej))
t("second call", lambda: rm.edge_lengths_device(decoder, z, ei, ej))
t("third call", lambda: rm.edge_lengths_device(decoder, z, ei, ej))
