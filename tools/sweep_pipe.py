"""Sweep of the streamed slot-major SSSP launch (csrc/sssp.cu: S sources through `slots` live tables, staggered
seeding) on the real C2 geodesic graph: slots x stagger, against lock-step batches of 16 (the round-1 flow).
usage: sweep_pipe.py [--slots 16,24,32] [--stagger 0,8,12,16] [--sources 64] [--trace]"""
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
ap.add_argument("--slots", default="16,24,32")
ap.add_argument("--stagger", default="0,8,12,16")
ap.add_argument("--sources", type=int, default=64)
ap.add_argument("--images", type=int, default=60000)
ap.add_argument("--trace", action="store_true")
a = ap.parse_args()

dev = torch.device("cuda")
z4 = torch.randn(a.images, 16, 7, 7, generator=torch.Generator().manual_seed(42))
torch.manual_seed(0)
decoder = SpatialVAE(**bench.FASHION_CFG).decoder.to(dev)
res = build_codebook_device(z4, decoder, k=20, sym="union", K=512, init="kpp", seed=42, batch_size=512, verbose=False)
g = res.g_lcc
med = np.asarray(res.medoids, dtype=np.int64)
del res, z4
S = a.sources
src = med[:S]
n, nnz = g.n, g.nnz
peak = 6552.0
try:
    peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    pass
D = torch.empty((S, n), dtype=torch.float32, device=dev)
Dref = torch.empty((S, n), dtype=torch.float32, device=dev)


def timed(fn, reps=4):
    ts = []
    st = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st = fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts[1:]), st


def lockstep():
    tot = None
    for b0 in range(0, S, 16):
        _, st = dg.sssp_multi(g, src[b0:b0 + 16], slots=16, out=Dref[b0:b0 + 16], stats=True, kernel="slot", stagger=0)
        tot = st if tot is None else {k: (tot[k] + v if k in ("relaxations", "rounds", "scans") else v) for k, v in st.items()}
    return tot


def trace(slots):
    total = _lib.load().geocb_sssp_workspace_bytes(n, slots)
    buf = [v_ for k_, v_ in dg._ws_cache.items() if k_[0] == "sssp"][0]
    tr = buf[total - 2048:total].view(torch.int64).cpu().numpy()
    nr = min(int(tr[7]), 120)
    t, q = tr[8:8 + 2 * nr:2], tr[9:9 + 2 * nr:2]
    ts = np.r_[t, tr[6]]
    print("   round: popped pairs -> us   " + "  ".join(f"{i}:{int(q[i])}->{(ts[i+1]-ts[i])/1e3:.0f}" for i in range(nr)), flush=True)


print(f"n={n} nnz={nnz} sources={S} peak={peak}", flush=True)
ms, st = timed(lockstep)
print(f"lock-step batches of 16: {ms:.2f} ms  {S*nnz/ms/1e6:.1f} G edge-relax/s  roofline {S*16*(nnz+n)/ms/1e6/peak:.3f} "
      f"rounds={st['rounds']} relax/edge={st['relaxations']/(S*nnz):.3f}", flush=True)
for slots in [int(x) for x in a.slots.split(",")]:
    for stg in [int(x) for x in a.stagger.split(",")]:
        ms, st = timed(lambda: dg.sssp_multi(g, src, slots=slots, out=D, stats=True, kernel="slot", stagger=stg)[1])
        print(f"streamed slots={slots} stagger={stg}: {ms:.2f} ms  {S*nnz/ms/1e6:.1f} G edge-relax/s  roofline "
              f"{S*16*(nnz+n)/ms/1e6/peak:.3f}  rounds={st['rounds']} scans={st['scans']} rounds/source={st['rounds_per_source']:.1f} "
              f"relax/edge={st['relaxations']/(S*nnz):.3f} same={bool(torch.equal(D, Dref))}", flush=True)
        if a.trace:
            trace(slots)
# the adaptive default (stagger from the measured rounds per source)
ms, st = timed(lambda: dg.sssp_multi(g, src, out=D, stats=True, kernel="slot")[1])
print(f"default (slots={st['slots']} stagger={st['stagger']}): {ms:.2f} ms  {S*nnz/ms/1e6:.1f} G edge-relax/s  roofline "
      f"{S*16*(nnz+n)/ms/1e6/peak:.3f} same={bool(torch.equal(D, Dref))}", flush=True)
