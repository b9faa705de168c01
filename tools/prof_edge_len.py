"""Per-kernel breakdown of the Riemannian edge-length stage (analysis only, not a bench number).
usage: prof_edge_len.py [EDGES]"""
import sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench
from vqvae_b200.models import SpatialVAE
from vqvae_b200.geo.riemannian_metric import edge_lengths_device

E = int(sys.argv[1]) if len(sys.argv) > 1 else 64 * 512 * 40
dev = torch.device("cuda")
torch.manual_seed(0)
decoder = SpatialVAE(**bench.FASHION_CFG).decoder.to(dev)
g = torch.Generator(device="cpu").manual_seed(1)
z = torch.randn(500000, 16, generator=g).to(dev)
ei = torch.randint(0, 500000, (E,), generator=g, dtype=torch.int32).to(dev)
ej = torch.randint(0, 500000, (E,), generator=g, dtype=torch.int32).to(dev)
edge_lengths_device(decoder, z, ei[:65536], ej[:65536]); torch.cuda.synchronize()
t0 = time.perf_counter(); out = edge_lengths_device(decoder, z, ei, ej); torch.cuda.synchronize(); dt = time.perf_counter() - t0
calls = E / (64 * 512)
print(f"E={E}: {dt*1e3:.1f} ms total, {dt*1e3/calls:.3f} ms per 32768-edge call, checksum {float(out.double().sum()):.6f}")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    edge_lengths_device(decoder, z, ei[:64 * 512 * 8], ej[:64 * 512 * 8]); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=70))
