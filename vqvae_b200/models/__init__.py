"""Decoder architectures the codebook path needs (reference: src/models/)."""
from .spatial_vae import SpatialVAE, SpatialDecoder, SpatialEncoder

__all__ = ["SpatialVAE", "SpatialDecoder", "SpatialEncoder"]
