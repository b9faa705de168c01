"""ReLU-kink margins of Riemannian edge lengths -- TEST / ANALYSIS INFRASTRUCTURE (imports only torch).

The edge length is built from JVPs of a ReLU network (reference: src/geo/riemannian_metric.py:12-35).  A JVP through
ReLU is discontinuous where a pre-activation crosses 0: if float32 rounding decides the sign of a pre-activation that
is ~1e-7 in magnitude, the edge length jumps by a finite amount, so two correct float32 implementations (different
summation orders) cannot agree to 1e-5 on such an edge.  This module measures, in float64, how close each edge comes
to a kink:

    margin(e) = min over both end-point decoder passes, every ReLU of the decoder and every unit of that ReLU of
                |pre-activation| of edge e's row

together with the float64 edge length.  The decoder is evaluated in the mode the caller left it in; in train mode
BatchNorm couples all rows of a `batch_size`-edge batch (build_codebook.py:27-29), so whole reference batches are
evaluated and `batch_margin` (the minimum over the batch) is returned too: one flipped unit in a batch shifts that
channel's tangent mean for every row of the batch by 1/(rows*H*W) of the jump.

The parity rule used by tests/test_gpu_pipeline.py::check_edge_lengths:
    every edge must agree with the reference to 1e-5 relative, except
      * edges with margin < EPS (their own ReLU mask is undecidable in float32), and
      * edges of a batch with batch_margin < EPS, which may deviate by at most COUPLED_TOL through the BatchNorm means.

CLI (measurement that justified EPS; output committed as profiles/r2_kink_margins.txt):
    python tools/kink_margin.py        # s1 case: reference float32 vs the same decoder in float64, on the CPU
"""
from __future__ import annotations

import copy

import numpy as np
import torch
import torch.nn as nn

EPS = 1e-5          # |pre-activation| below which a ReLU mask is treated as undecidable (activations are O(1)).  Largest margin seen
                    # on a flipped edge: 5.4e-7 with fp32 SIMT GEMMs (reference fp32 vs fp64, and GEOCB_JVP_GEMM=cublas), 4.2e-6 with the
                    # tcgen05 3xTF32 GEMM, whose error is 1.3-1.9e-6 x sum|a||w| against 3e-7 (tests/test_gpu_jvp_gemm.py)
COUPLED_TOL = 1e-4  # bound on the deviation a flipped unit elsewhere in the BatchNorm batch can cause


@torch.no_grad()
def relu_margins(decoder, z_start, z_end, batch_size, edges):
    """float64 margins for the listed edge positions.  Returns dict(edge -> (margin, batch_margin, length64))."""
    dec = copy.deepcopy(decoder).double().cpu()
    flat_in = hasattr(next(dec.children()), "in_features")
    zs, ze = z_start.double().cpu(), z_end.double().cpu()
    relus = [m for m in dec.modules() if isinstance(m, nn.ReLU)]
    captured = []
    hooks = [m.register_forward_pre_hook(lambda mod, inp: captured.append(inp[0].detach().clone())) for m in relus]
    out = {}
    try:
        by_batch = {}
        for e in np.asarray(edges, dtype=np.int64).tolist():
            by_batch.setdefault(e // batch_size, []).append(e)
        for b, es in by_batch.items():
            lo, hi = b * batch_size, min((b + 1) * batch_size, zs.shape[0])
            a, c = zs[lo:hi], ze[lo:hi]
            d = c - a
            row_margin = torch.full((hi - lo,), float("inf"), dtype=torch.float64)
            norms = []
            for base in (a, c):
                captured.clear()

                def img(p):
                    if not flat_in and p.ndim == 2:
                        p = p[:, :, None, None]
                    y = torch.sigmoid(dec(p))
                    return y.reshape(y.shape[0], -1)

                with torch.enable_grad():
                    j = torch.autograd.functional.jvp(img, (base.clone().requires_grad_(True),), (d,))[1]
                norms.append(torch.linalg.vector_norm(j, dim=1))
                # (functional.jvp runs the forward more than once; every captured pass is the same primal)
                for x in captured:
                    row_margin = torch.minimum(row_margin, x.reshape(x.shape[0], -1).abs().min(dim=1).values)
            L64 = 0.5 * (norms[0] + norms[1])
            bm = float(row_margin.min())
            for e in es:
                out[e] = (float(row_margin[e - lo]), bm, float(L64[e - lo]))
    finally:
        for h in hooks:
            h.remove()
    return out


def check_edge_lengths(got, ref, decoder, z_start, z_end, batch_size, positions=None, tol=1e-5):
    """The per-edge rule above.  `got`/`ref` are the lengths of the edges at `positions` (default: all edges, in
    order) of the (z_start, z_end) edge list.  Returns a report dict; raises AssertionError on a violation."""
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    pos = np.arange(len(ref)) if positions is None else np.asarray(positions, dtype=np.int64)
    rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-30)
    bad = np.nonzero(rel > tol)[0]
    report = {"edges": int(len(ref)), "beyond_tol": int(len(bad)), "max_rel": float(rel.max()) if len(rel) else 0.0,
              "median_rel": float(np.median(rel)) if len(rel) else 0.0, "own_kink": 0, "batch_kink": 0}
    assert report["median_rel"] <= 1e-6, report
    assert len(bad) <= max(8, 0.002 * len(ref)), f"{len(bad)} of {len(ref)} edges beyond {tol}: not a kink effect"
    if len(bad):
        marg = relu_margins(decoder, z_start, z_end, batch_size, pos[bad])
        for i in bad:
            m, bm, _ = marg[int(pos[i])]
            if m < EPS:
                report["own_kink"] += 1
            elif bm < EPS and rel[i] <= COUPLED_TOL:
                report["batch_kink"] += 1
            else:
                raise AssertionError(f"edge {int(pos[i])}: rel err {rel[i]:.3g} with ReLU margin {m:.3g} "
                                     f"(batch margin {bm:.3g}) -- not explained by a ReLU kink")
    return report


if __name__ == "__main__":
    import sys
    from pathlib import Path
    ROOT = Path(__file__).resolve().parents[1]
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "oracle"))
    sys.path.insert(0, str(ROOT / "tests"))
    import geo_oracle as oracle
    from conftest import s1_latents, fashion_decoder
    s1 = dict(np.load(ROOT / "tests" / "golden" / "s1_stages.npz"))
    _, zf = s1_latents()
    dec = fashion_decoder()
    a, b = torch.from_numpy(zf[s1["ei"]]), torch.from_numpy(zf[s1["ej"]])
    L32 = oracle.edge_lengths(dec, a, b, 512).numpy()
    assert np.array_equal(L32, s1["edge_len"]), "oracle float32 lengths differ from the golden vector"
    E = len(L32)
    marg = relu_margins(dec, a, b, 512, np.arange(E))
    m = np.array([marg[e][0] for e in range(E)])
    bm = np.array([marg[e][1] for e in range(E)])
    L64 = np.array([marg[e][2] for e in range(E)])
    rel = np.abs(L32 - L64) / L64
    print(f"s1: {E} edges; reference float32 vs float64 of the same decoder (train-mode BatchNorm, batch 512)")
    print(f"  rel err: median {np.median(rel):.3g}, >1e-5: {(rel > 1e-5).sum()}, >1e-4: {(rel > 1e-4).sum()}, max {rel.max():.3g}")
    for eps in (1e-7, 1e-6, 1e-5, 1e-4):
        print(f"  margin < {eps:g}: {(m < eps).sum()} edges ({(bm < eps).sum()} edges sit in a batch with such a unit)")
    big = np.nonzero(rel > 1e-5)[0]
    print("  edges beyond 1e-5: (edge, rel err, own margin, batch margin)")
    for e in big:
        print(f"    {e:6d}  {rel[e]:.3g}  {m[e]:.3g}  {bm[e]:.3g}")
    unexplained = [e for e in big if not (m[e] < EPS or (bm[e] < EPS and rel[e] <= COUPLED_TOL))]
    print(f"  unexplained by the rule (EPS={EPS:g}, COUPLED_TOL={COUPLED_TOL:g}): {len(unexplained)}")
