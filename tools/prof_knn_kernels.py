"""Kernel-level split of geocb_knn_topk at C2 size (torch profiler, CUDA activities only)."""
import sys
from pathlib import Path
import torch
from torch.profiler import profile, ProfilerActivity
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg
images = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 16
z = torch.randn(images * 49, d, generator=torch.Generator().manual_seed(42)).cuda()
dg.knn_topk(z, 21)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    dg.knn_topk(z, 21)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=12, max_name_column_width=90))
