from vqvae_b200.models import SpatialVAE

__all__ = ["SpatialVAE"]
