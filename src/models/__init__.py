"""`src.models`: SpatialVAE is this repo's (state-dict compatible with the reference's, spatial_vae.py:47-81); the
reference's other models (`src.models.vae.VAE` used by the legacy drivers, `transformer`) resolve in its checkout
when one is reachable (see src/__init__.py)."""
from vqvae_b200.models import SpatialVAE

from .. import overlay as _overlay

_overlay(__path__, "models")

__all__ = ["SpatialVAE"]


def __getattr__(name):  # `from src.models import VAE` / `Transformer` as in the reference's package init
    if name in ("VAE", "Transformer"):
        import importlib
        mod = importlib.import_module(f"{__name__}.{'vae' if name == 'VAE' else 'transformer'}")
        return getattr(mod, name)
    raise AttributeError(name)
