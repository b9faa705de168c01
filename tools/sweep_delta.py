"""Bucket-width / slot / library-variant sweep of the SSSP on the real C2 geodesic graph (built once, cached under /tmp).
usage: sweep_delta.py [SCALES] [SLOTS] [LIB.so ...]    e.g.  sweep_delta.py 0.5,1,2,4 16
Extra arguments are alternative builds of libgeocb.so (compile-time knobs of sssp.cu); each is loaded in turn in the
same process, after the default library."""
import os, sys, time
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench
from vqvae_b200 import _lib, device_graph as dg
from vqvae_b200.models import SpatialVAE
from vqvae_b200.pipeline import build_codebook_device

scales = [float(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0.5,1,2,4").split(",")]
slot_list = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "16").split(",")]
dev = torch.device("cuda")
cache = Path("/tmp/c2_geo_graph.pt")
if cache.exists():
    ip, ix, dt, med = torch.load(cache)
    g = dg.DeviceCSR(ip.to(dev), ix.to(dev), dt.to(dev), ip.numel() - 1)
else:
    z4 = torch.randn(60000, 16, 7, 7, generator=torch.Generator().manual_seed(42))
    torch.manual_seed(0)
    decoder = SpatialVAE(**bench.FASHION_CFG).decoder.to(dev)
    res = build_codebook_device(z4, decoder, k=20, sym="union", K=512, init="kpp", seed=42, batch_size=512, verbose=False)
    g = res.g_lcc
    med = torch.as_tensor(np.asarray(res.medoids, dtype=np.int64))
    torch.save((g.indptr.cpu(), g.indices.cpu(), g.data.cpu(), med), cache)
src = med.numpy()[:64]
_, _, mean_w = g.adjacency(False)
D = torch.empty((64, g.n), dtype=torch.float32, device=dev)
ref = None
libs = [None] + [a for a in sys.argv[3:]]
import ctypes
from pathlib import Path as _P
for lib_path in libs:
  if lib_path is not None:
    _lib._lib = None
    _lib.LIB_PATH = _P(lib_path)
    _lib.load()
    print(f"--- {lib_path}", flush=True)
  for slots in slot_list:
    for sc in scales:
          ts = []
          for it in range(3):
              e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
              e0.record()
              _, st = dg.sssp_multi(g, src, slots=slots, out=D, delta=mean_w * sc, stats=True)
              e1.record(); torch.cuda.synchronize()
              ts.append(e0.elapsed_time(e1))
          ms = min(ts[1:])
          chk = float(D[:4].double().sum())
          ref = chk if ref is None else ref
          print(f"slots={slots} delta={sc:g}x mean: {ms:.2f} ms  {64*g.nnz/ms/1e6:.1f} G edge-relax/s  rounds={st['rounds']} scans={st['scans']} "
                f"relax/edge={st['relaxations']/(64*g.nnz):.3f} same={chk == ref}", flush=True)
