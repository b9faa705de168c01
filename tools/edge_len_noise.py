"""How many edges leave the 1e-5 band, per GEMM back end (GEOCB_JVP_GEMM=tc|cublas), on the s1 golden case.
Prints count beyond 1e-5 / 1e-4 and the largest ReLU margin among them (tools/kink_margin.py)."""
import os, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "tests", ROOT / "tools", ROOT / "oracle"):
    sys.path.insert(0, str(p))
from conftest import s1_latents, fashion_decoder
import kink_margin
from vqvae_b200.geo.riemannian_metric import edge_lengths_riemannian
s1 = dict(np.load(ROOT / "tests" / "golden" / "s1_stages.npz"))
_, zf = s1_latents()
a, b = torch.from_numpy(zf[s1["ei"]]), torch.from_numpy(zf[s1["ej"]])
for mode in ("cublas", "tc"):
    os.environ["GEOCB_JVP_GEMM"] = mode
    L = edge_lengths_riemannian(fashion_decoder().cuda(), a, b, 512).cpu().numpy()
    rel = np.abs(L - s1["edge_len"]) / s1["edge_len"]
    bad = np.nonzero(rel > 1e-5)[0]
    marg = kink_margin.relu_margins(fashion_decoder(), a, b, 512, bad) if len(bad) else {}
    own = sorted(marg[int(e)][0] for e in bad)
    print(f"{mode}: median {np.median(rel):.3g}  >1e-5: {len(bad)}  >1e-4: {(rel > 1e-4).sum()}  max {rel.max():.3g}  "
          f"margins of the offenders: {['%.2g' % m for m in own]}", flush=True)
