"""Time geocb_knn_topk (+ symmetrise) on synthetic latents: prof_knn.py IMAGES [d]  (latents = IMAGES*49)."""
import sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg
images = int(sys.argv[1]); d = int(sys.argv[2]) if len(sys.argv) > 2 else 16
dev = torch.device("cuda")
z = torch.randn(images * 49, d, generator=torch.Generator().manual_seed(42)).to(dev)
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    idx, dist, nun = dg.knn_topk(z, 21)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    M = z.shape[0]
    print(f"M={M} d={d} knn_topk {t1-t0:.3f}s  {M*M/(t1-t0)/1e12:.2f} T pair-scores/s  exact-fallback rows={int(nun)}  "
          f"self-first={bool((idx[:,0]==torch.arange(M,device=dev,dtype=torch.int32)).all())}", flush=True)
