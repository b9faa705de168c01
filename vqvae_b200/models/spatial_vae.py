"""SpatialVAE architecture -- parameter-compatible with the reference's
`src/models/spatial_vae.py` (state_dict keys, shapes and construction order match, so a
reference checkpoint loads and `torch.manual_seed(s); SpatialVAE(**cfg)` yields the same
random init).  Only the decoder is used by the codebook path (build_codebook.py:27-29);
the encoder exists so checkpoints load with strict=True.  Training/loss code is out of scope.
"""
from __future__ import annotations

from typing import Sequence

import torch
import torch.nn as nn


def make_norm(norm_type: str, channels: int) -> nn.Module:
    """batch | group | anything else -> identity (spatial_vae.py:8-19)."""
    kind = (norm_type or "none").lower()
    if kind == "batch":
        return nn.BatchNorm2d(channels)
    if kind == "group":
        g = max(1, min(32, channels))
        while g > 1 and channels % g:
            g -= 1
        return nn.GroupNorm(g, channels)
    return nn.Identity()


class SpatialEncoder(nn.Module):
    """stride-2 3x3 conv stack + 1x1 heads (spatial_vae.py:22-44)."""

    def __init__(self, input_channels: int, channels: Sequence[int], latent_dim: int, norm_type: str):
        super().__init__()
        blocks, c_in = [], input_channels
        for c in channels:
            blocks += [nn.Conv2d(c_in, c, 3, stride=2, padding=1), make_norm(norm_type, c), nn.ReLU(inplace=True)]
            c_in = c
        self.conv_layers = nn.Sequential(*blocks)
        self.fc_mu = nn.Conv2d(c_in, latent_dim, 1)
        self.fc_logvar = nn.Conv2d(c_in, latent_dim, 1)

    def forward(self, x):
        h = self.conv_layers(x)
        return self.fc_mu(h), self.fc_logvar(h)


class SpatialDecoder(nn.Module):
    """1x1 conv -> two (ConvT k4 s2 p1 + norm + ReLU) -> ConvT k4 s2 (p=1 for 32px, p=3 for 28px)
    (spatial_vae.py:47-81)."""

    def __init__(self, out_channels: int, channels: Sequence[int], latent_dim: int, output_image_size: int,
                 norm_type: str):
        super().__init__()
        if output_image_size not in (28, 32):
            raise ValueError(f"Unsupported output size: {output_image_size}")
        self.conv_in = nn.Conv2d(latent_dim, channels[0], 1)
        last_pad = 1 if output_image_size == 32 else 3
        self.deconv_layers = nn.Sequential(
            nn.ConvTranspose2d(channels[0], channels[1], 4, stride=2, padding=1),
            make_norm(norm_type, channels[1]),
            nn.ReLU(inplace=True),
            nn.ConvTranspose2d(channels[1], channels[2], 4, stride=2, padding=1),
            make_norm(norm_type, channels[2]),
            nn.ReLU(inplace=True),
            nn.ConvTranspose2d(channels[2], out_channels, 4, stride=2, padding=last_pad),
        )

    def forward(self, z):
        return self.deconv_layers(self.conv_in(z))


class SpatialVAE(nn.Module):
    """Container with the reference's attribute names (spatial_vae.py:84-125)."""

    def __init__(self, in_channels, enc_channels, dec_channels, latent_dim, recon_loss, output_image_size,
                 norm_type, **kwargs):
        super().__init__()
        self.encoder = SpatialEncoder(in_channels, enc_channels, latent_dim, norm_type)
        self.decoder = SpatialDecoder(in_channels, dec_channels, latent_dim, output_image_size, norm_type)
        assert recon_loss in {"bce", "mse"}
        self.recon_loss = recon_loss
        self.mse_use_sigmoid = kwargs.get("mse_use_sigmoid", True)

    def forward(self, x):
        mu, logvar = self.encoder(x)
        z = mu + torch.randn_like(mu) * torch.exp(0.5 * logvar)
        return self.decoder(z), mu, logvar, z
