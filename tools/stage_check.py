"""Run one stage of the hot path on tiny inputs and print progress (used under `timeout` on the GPU box)."""
import sys, time
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests" / "golden")]
from vqvae_b200 import device_graph as dg
import geo_oracle as oracle

stage = sys.argv[1]
dev = torch.device("cuda")
z = torch.randn(2000, 16, generator=torch.Generator().manual_seed(1))
def P(*a): print(*a, flush=True)
t0 = time.time()
if stage == "knn":
    idx, dist, nun = dg.knn_topk(z.to(dev), 11); torch.cuda.synchronize()
    P("knn done", time.time() - t0, "exact rows", int(nun))
    Wo, io = oracle.knn_graph(z.numpy(), k=10, mode="distance", sym="union")
    P("idx equal", np.array_equal(idx[:, 1:].cpu().numpy(), io["indices"]))
    g = dg.symmetrize(idx[:, 1:].contiguous(), dist[:, 1:].contiguous(), "union"); torch.cuda.synchronize()
    W = g.to_scipy()
    P("csr equal", np.array_equal(W.indptr, Wo.indptr), np.array_equal(W.indices, Wo.indices), np.abs(W.data - Wo.data).max())
    ei, ej, pij, pji = dg.upper_edges(g); torch.cuda.synchronize(); P("upper", ei.numel())
    m, nid, nk, nc = dg.largest_component(g); P("lcc", nk, nc)
elif stage in ("sssp", "voronoi", "kpp"):
    Wo, _ = oracle.knn_graph(z.numpy(), k=10, mode="distance", sym="union")
    g = dg.DeviceCSR.from_scipy(Wo, dev)
    g.adjacency(); torch.cuda.synchronize(); P("adj built", g.adjacency()[2])
    if stage == "sssp":
        for slots in (1, 3):
            D, st = dg.sssp_multi(g, [0, 17, 1999], slots=slots, stats=True)
            P("sssp slots", slots, st, "bit-equal", np.array_equal(D.cpu().numpy(), oracle.sssp_multi(Wo, [0, 17, 1999])))
        D, Pp = dg.sssp_multi(g, [5], want_pred=True); P("pred ok", int((Pp < 0).sum()))
    elif stage == "voronoi":
        st = dg.VoronoiState(g)
        for c in (3, 500, 1500, 3):
            st.insert(c); torch.cuda.synchronize(); P("inserted", c)
        D = oracle.sssp_multi(Wo, [3, 500, 1500])
        P("dmin equal", np.array_equal(st.dmin32.cpu().numpy(), D.min(0)), "label equal", np.array_equal(st.label.cpu().numpy(), D.argmin(0)))
    else:
        from vqvae_b200.geo.kmeans_optimized import fit_kmedoids_optimized
        got = fit_kmedoids_optimized(Wo, K=16, seed=42); exp = oracle.fit_kmedoids(Wo, K=16, seed=42)
        P("kmedoids equal", np.array_equal(got[0], exp[0]), np.array_equal(got[1], exp[1]), got[2], exp[2])
P(stage, "finished in", time.time() - t0)
