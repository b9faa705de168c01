"""SSSP-only driver for timing / ncu: kNN graph of synthetic latents with hash weights (no JVP stage).
usage: prof_sssp.py IMAGES SLOTS[,SLOTS..] [LAUNCHES] [--trace] [--sources S]   (latents = IMAGES*49)
The graph is cached under /tmp so several library builds can be compared in one gpurun call."""
import os, sys, time
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import _lib, device_graph as dg
from vqvae_b200.geo.knn_graph_optimized import knn_graph_device

S_total = 0
argv = sys.argv[1:]
if "--sources" in argv:  # stream this many sources through the slots (default: sources = slots)
    i = argv.index("--sources"); S_total = int(argv[i + 1]); del argv[i:i + 2]
args = [a for a in argv if not a.startswith("--")]
images, slot_list = int(args[0]), [int(x) for x in args[1].split(",")]
launches = int(args[2]) if len(args) > 2 else 3
trace = "--trace" in sys.argv
dev = torch.device("cuda")
cache = Path(f"/tmp/prof_graph_{images}.pt")
t0 = time.time()
if cache.exists():
    ip, ix, dt = torch.load(cache)
    g = dg.DeviceCSR(ip.to(dev), ix.to(dev), dt.to(dev), ip.numel() - 1)
else:
    z = torch.randn(images, 16, 7, 7, generator=torch.Generator().manual_seed(42)).permute(0, 2, 3, 1).reshape(-1, 16).contiguous().to(dev)
    g, _, _ = knn_graph_device(z, 20, "connectivity", "union")
    # symmetric pseudo-random weights in [0.25, 2.25): integer hash of (min(i,j), max(i,j))
    rows = torch.repeat_interleave(torch.arange(g.n, device=dev), g.indptr[1:] - g.indptr[:-1])
    cols = g.indices.long()
    a, b = torch.minimum(rows, cols), torch.maximum(rows, cols)
    h = (a * 2654435761 + b * 40503) & 0xFFFFFFFF
    h = ((h ^ (h >> 15)) * 2246822519) & 0xFFFFFFFF
    h = (h ^ (h >> 13)) & 0xFFFFF
    g.data = (0.25 + 2.0 * h.double() / (1 << 20)).float()
    del rows, cols, a, b, h
    torch.save((g.indptr.cpu(), g.indices.cpu(), g.data.cpu()), cache)
torch.cuda.synchronize(); t_graph = time.time() - t0
g.adjacency(); torch.cuda.synchronize()
tag = os.environ.get("PROF_TAG", "")
for slots in slot_list:
    rng = np.random.RandomState(0)
    times = []
    for i in range(launches):
        S = S_total or slots
        src = rng.choice(g.n, S, replace=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        try:
            D, st = dg.sssp_multi(g, src, slots=slots, stats=True)
        except RuntimeError as ex:
            print("FAILED:", ex, "sources", src.tolist(), flush=True)
            break
        e1.record(); torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    if not times:
        continue
    ms = float(np.median(times[1:] or times))
    print(f"{tag} n={g.n} nnz={g.nnz} slots={slots} sources={S}: {ms:.3f} ms  {S*g.nnz/ms/1e6:.2f} G edge-relax/s  "
          f"roofline {S*16*(g.nnz+g.n)/ms/1e6/6538.9:.3f}  rounds={st['rounds']} scans={st['scans']} "
          f"relax/edge={st['relaxations']/(S*g.nnz):.3f} graph_s={t_graph:.1f}", flush=True)
    if trace:
        # round trace of the last launch (written by the kernel into the tail of its workspace)
        total = _lib.load().geocb_sssp_workspace_bytes(g.n, slots)
        phases = os.environ.get("PROF_PHASES") == "1"  # library built with -DGEOCB_SSSP_TRACE_PHASES (4096-byte stats)
        sb = 4096 if phases else 2048
        tr = [v for k, v in dg._ws_cache.items() if k[0] == "sssp"][0][total - sb:total].view(torch.int64).cpu().numpy()
        nr = int(tr[7]); t = tr[8:8 + 2 * min(nr, 120):2]; q = tr[9:9 + 2 * min(nr, 120):2]
        ts = np.r_[t, tr[6]]
        print(f"improvements/relax={st['improvements']/max(st['relaxations'],1):.4f}")
        if phases:
            nrr = min(nr, 120)
            a_end, b_end = tr[256:256 + nrr], tr[384:384 + nrr]
            print("round: sweep us + control/scan us   " + "  ".join(
                f"{i}:{(a_end[i]-t[i])/1e3:.0f}+{(b_end[i]-a_end[i])/1e3:.0f}" for i in range(nrr)), flush=True)
        print("round: popped -> us   " + "  ".join(f"{i}:{int(q[i])}->{(ts[i+1]-ts[i])/1e3:.0f}" for i in range(len(t))), flush=True)
