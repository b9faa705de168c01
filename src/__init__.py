"""Import shim: lets reference-side code (`from src.geo... import ...`, `python src/scripts/build_codebook.py`) run
against vqvae_b200 unchanged.

Overlay: when a checkout of the reference is reachable (env `VQVAE_REFERENCE_ROOT`, or a later `sys.path` entry that
holds the reference's `src/` package), its directories are appended to this package's search path.  Modules this
repo implements (`src.geo.*`, `src.scripts.build_codebook`, `src.models.SpatialVAE`) win; everything else
(`src.training.*` incl. the legacy YAML drivers, `src.utils`, `src.data`, `src.eval`, `src.models.vae`, ...) resolves
in the reference, so its drivers, demos and tests run with the geodesic path on the GPU (level 1)."""
import os
import sys
from pathlib import Path

_HERE = Path(__file__).resolve().parent


def reference_src():
    """Path of the reference's `src/` directory, or None."""
    cands = []
    env = os.environ.get("VQVAE_REFERENCE_ROOT")
    if env:
        cands.append(Path(env) / "src")
    for p in sys.path:
        try:
            cands.append(Path(p or ".").resolve() / "src")
        except OSError:
            continue
    for c in cands:
        try:
            if c.resolve() != _HERE and (c / "geo" / "knn_graph_optimized.py").exists() and (c / "__init__.py").exists():
                return c
        except OSError:
            continue
    return None


def overlay(pkg_path, sub=""):
    """Append the reference's matching directory to a shim package's __path__ (idempotent)."""
    ref = reference_src()
    if ref is None:
        return None
    d = ref / sub if sub else ref
    if d.is_dir() and str(d) not in pkg_path:
        pkg_path.append(str(d))
    return d


overlay(__path__)

from vqvae_b200.geo import build_knn_graph, dijkstra_multi_source  # noqa: E402

__all__ = ["build_knn_graph", "dijkstra_multi_source"]
