"""Same path as the reference's src/models/vae.py; the inference-side VAE of vqvae_b200 (checkpoint compatible)."""
from vqvae_b200.models.vae import VAE, Decoder, Encoder  # noqa: F401
