"""Per-phase clock totals of the tcgen05 kNN filter (build: make -C vqvae_b200/csrc knntrace; clock64 differences
accumulated in registers, one atomicAdd per warp at the end).  usage: prof_knn_trace.py IMAGES [d]"""
import ctypes as C
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import _lib
_lib.LIB_PATH = ROOT / "vqvae_b200" / "libgeocb_knntrace.so"
from vqvae_b200 import device_graph as dg
images = int(sys.argv[1]); d = int(sys.argv[2]) if len(sys.argv) > 2 else 16
z = torch.randn(images * 49, d, generator=torch.Generator().manual_seed(42)).cuda()
lib = _lib.load()
out = (C.c_ulonglong * 16)()
dg.knn_topk(z, 21)
torch.cuda.synchronize()
lib.geocb_knn_trace_read(out, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); dg.knn_topk(z, 21); e1.record(); torch.cuda.synchronize()
lib.geocb_knn_trace_read(out, 1)
t = list(out)
n, m = max(1, t[5]), max(1, t[11])
tot = t[0] + t[2] + t[3]
print(f"M={z.shape[0]} d={d}: knn_topk {e0.elapsed_time(e1):.1f} ms (instrumented build); {t[11]} CTA-chunks, {t[5]} epilogue warp-chunks")
for nm, val in (("wait for the accumulator", t[0]), ("TMEM loads", t[2]), ("screen + insert", t[3])):
    print(f"  epilogue warp, {nm:28s} {val/n:8.1f} clk per chunk  ({100*val/tot:5.1f} %)")
print(f"  epilogue warp, total                        {tot/n:8.1f} clk per chunk")
for nm, val in (("wait for the B stage (TMA)", t[8]), ("wait for drained accumulators", t[9]), ("issue MMAs + commits", t[10] - t[9])):
    print(f"  MMA issuer,    {nm:28s} {val/m:8.1f} clk per chunk")
