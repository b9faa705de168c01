"""Decoder architectures the codebook path needs (reference: src/models/)."""
from .spatial_vae import SpatialVAE, SpatialDecoder, SpatialEncoder
from .vae import VAE, Decoder, Encoder

__all__ = ["SpatialVAE", "SpatialDecoder", "SpatialEncoder", "VAE", "Decoder", "Encoder"]
