"""Sweep of the vertex-major SSSP kernel (csrc/sssp_vm.cu) against the slot-major one on the real C2 geodesic graph
(built once per box, cached under /tmp):  destination blocks B, bucket width, unblocked-round threshold.
usage: sweep_vm.py [--blocks 1,8,15,24] [--scales 2,4,8] [--minact 16] [--trace] [--sources 64] [--images 60000]
Prints ms per 64-source step, G edge-relax/s (S*nnz/t), roofline fraction, rounds, relaxations and edge visits per
algorithmic edge, and whether the distances are bit-identical to the slot-major kernel's."""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from vqvae_b200 import _lib, device_graph as dg  # noqa: E402
from vqvae_b200.models import SpatialVAE  # noqa: E402
from vqvae_b200.pipeline import build_codebook_device  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--blocks", default="1,8,15,24")
ap.add_argument("--scales", default="4")
ap.add_argument("--minact", default="16", help="unblocked below n/X popped vertices")
ap.add_argument("--sources", type=int, default=64)
ap.add_argument("--images", type=int, default=60000)
ap.add_argument("--trace", action="store_true")
ap.add_argument("--skip-slot", action="store_true")
ap.add_argument("--gen", default="16", help="vertex-major generations to sweep: 16 (sssp_vm.cu), 32 (sssp_vm32.cu)")
ap.add_argument("--libs", default="", help="comma-separated alternative builds of libgeocb.so, swept after the default one")
a = ap.parse_args()

dev = torch.device("cuda")
cache = Path(f"/tmp/c2_geo_graph_{a.images}.pt")
if cache.exists():
    ip, ix, dt, med = torch.load(cache)
    g = dg.DeviceCSR(ip.to(dev), ix.to(dev), dt.to(dev), ip.numel() - 1)
else:
    z4 = torch.randn(a.images, 16, 7, 7, generator=torch.Generator().manual_seed(42))
    torch.manual_seed(0)
    decoder = SpatialVAE(**bench.FASHION_CFG).decoder.to(dev)
    res = build_codebook_device(z4, decoder, k=20, sym="union", K=512, init="kpp", seed=42, batch_size=512, verbose=False)
    g = res.g_lcc
    med = torch.as_tensor(np.asarray(res.medoids, dtype=np.int64))
    torch.save((g.indptr.cpu(), g.indices.cpu(), g.data.cpu(), med), cache)
    del res, z4
S = a.sources
src = med.numpy()[:S]
_, _, mean_w = g.adjacency(False)
n, nnz = g.n, g.nnz
peak = 6552.0
try:
    peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    pass
D = torch.empty((S, n), dtype=torch.float32, device=dev)
Dref = torch.empty((S, n), dtype=torch.float32, device=dev)


def timed(fn, reps=3):
    ts = []
    st = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _, st = fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts[1:] or ts), st


def report(tag, ms, st, same):
    print(f"{tag}: {ms:.2f} ms  {S*nnz/ms/1e6:.1f} G edge-relax/s  roofline {S*16*(nnz+n)/ms/1e6/peak:.3f}  rounds={st['rounds']} "
          f"scans={st['scans']} relax/edge={st['relaxations']/(S*nnz):.3f} "
          f"visits/edge={st.get('edge_visits', 0)/(S*nnz):.3f} improv/relax={st['improvements']/max(1,st['relaxations']):.3f} "
          f"blocked_rounds={st.get('blocked_rounds', '-')} same={same}", flush=True)


print(f"n={n} nnz={nnz} mean_w={mean_w:.4f} sources={S} peak={peak}", flush=True)
if not a.skip_slot:
    ms, st = timed(lambda: dg.sssp_multi(g, src, slots=16, out=Dref, stats=True, kernel="slot"))
    report("slot-major 16 slots delta=4x", ms, st, True)
    if a.trace:  # round trace of the last slot-major launch (same stats layout, tail of its workspace)
        total = _lib.load().geocb_sssp_workspace_bytes(n, 16)
        tr = [v_ for k_, v_ in dg._ws_cache.items() if k_[0] == "sssp"][0][total - 2048:total].view(torch.int64).cpu().numpy()
        nr = min(int(tr[7]), 120)
        t, q = tr[8:8 + 2 * nr:2], tr[9:9 + 2 * nr:2]
        ts = np.r_[t, tr[6]]
        print("   last launch, round: popped (slot,vertex) pairs -> us   " +
              "  ".join(f"{i}:{int(q[i])}->{(ts[i+1]-ts[i])/1e3:.0f}" for i in range(nr)), flush=True)
        print("   relaxations/s by round (popped x mean degree / round time):  " +
              "  ".join(f"{i}:{q[i]*nnz/n/max(ts[i+1]-ts[i],1):.0f}G" for i in range(nr)), flush=True)
else:
    dg.sssp_multi(g, src, slots=16, out=Dref, kernel="slot")
for lib_path in [None] + [x for x in a.libs.split(",") if x]:
  if lib_path is not None:
    _lib._lib = None
    _lib.LIB_PATH = Path(lib_path)
    _lib.load()
    print(f"--- {lib_path}", flush=True)
  for gen in [int(x) for x in a.gen.split(",")]:
   wide = gen == 32
   for B in [int(x) for x in a.blocks.split(",")]:
    for sc in [float(x) for x in a.scales.split(",")]:
        for mf in [int(x) for x in a.minact.split(",")]:
            ms, st = timed(lambda: dg.sssp_multi_vm(g, src, out=D, stats=True, blocks=B, delta=sc * mean_w,
                                                    blk_min_active=max(1, n // mf), wide=wide))
            report(f"vm{gen} B={B} delta={sc:g}x minact=n/{mf}", ms, st, bool(torch.equal(D, Dref)))
            if a.trace:
              if True:
                    L = _lib.load()
                    total = L.geocb_sssp_vm32_workspace_bytes(n) if wide else L.geocb_sssp_vm_workspace_bytes(n)
                    key = "sssp_vm32" if wide else "sssp_vm"
                    tr = [v_ for k_, v_ in dg._ws_cache.items() if k_[0] == key][0][total - 2048:total].view(torch.int64).cpu().numpy()
                    nr = min(int(tr[7]), 120)
                    t, q = tr[8:8 + 2 * nr:2], tr[9:9 + 2 * nr:2]
                    ts = np.r_[t, tr[6]]
                    print("   last launch, round: popped (slot,vertex) pairs -> us   " +
                          "  ".join(f"{i}:{int(q[i])}->{(ts[i+1]-ts[i])/1e3:.0f}" for i in range(nr)), flush=True)
