"""Flat-latent ("vanilla") VAE of the reference (src/models/vae.py:22-110), inference side only: the architecture and
parameter names are kept so reference checkpoints load with strict=True (`encoder.conv_layers.*`, `encoder.fc_mu`,
`encoder.fc_logvar`, `decoder.fc`, `decoder.deconv1.*`, `decoder.deconv2.*`, `decoder.output_layer`); the training
losses are out of scope (only the decoder enters the codebook path, through its JVPs)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .spatial_vae import make_norm as _get_norm_layer


def _conv_block(cin, cout, norm_type):
    return [nn.Conv2d(cin, cout, 3, stride=2, padding=1), _get_norm_layer(norm_type, cout), nn.ReLU(inplace=True)]


class Encoder(nn.Module):
    """Three stride-2 convolutions -> flat (mu, logvar) (src/models/vae.py:22-50)."""

    def __init__(self, input_channels=1, channels=(32, 64, 128), latent_dim=16, norm_type="none"):
        super().__init__()
        blocks, cin = [], input_channels
        for c in channels:
            blocks += _conv_block(cin, c, norm_type)
            cin = c
        self.conv_layers = nn.Sequential(*blocks)
        self.feature_dim = channels[-1] * 16
        self.fc_mu = nn.Linear(self.feature_dim, latent_dim)
        self.fc_logvar = nn.Linear(self.feature_dim, latent_dim)

    def forward(self, x):
        h = self.conv_layers(x).flatten(1)
        return self.fc_mu(h), self.fc_logvar(h)


class Decoder(nn.Module):
    """Linear -> (C,4,4) -> three transposed convolutions (src/models/vae.py:53-85)."""

    def __init__(self, out_channels=1, channels=(128, 64, 32), latent_dim=16, output_image_size=28, norm_type="none"):
        super().__init__()
        c0, c1, c2 = channels
        self.fc = nn.Linear(latent_dim, c0 * 16)
        # 4x4 -> 8x8 (32-px images) or 7x7 (28-px images)
        self.deconv1 = nn.Sequential(
            nn.ConvTranspose2d(c0, c1, 3, stride=2, padding=1, output_padding=1 if output_image_size == 32 else 0),
            _get_norm_layer(norm_type, c1), nn.ReLU(inplace=True))
        self.deconv2 = nn.Sequential(nn.ConvTranspose2d(c1, c2, 4, stride=2, padding=1), _get_norm_layer(norm_type, c2),
                                     nn.ReLU(inplace=True))
        self.output_layer = nn.ConvTranspose2d(c2, out_channels, 4, stride=2, padding=1)

    def forward(self, z):
        h = self.fc(z)
        h = h.view(h.size(0), -1, 4, 4)
        return self.output_layer(self.deconv2(self.deconv1(h)))


class VAE(nn.Module):
    def __init__(self, in_channels=1, enc_channels=(32, 64, 128), dec_channels=(128, 64, 32), latent_dim=16,
                 recon_loss="bce", output_image_size=28, norm_type="none", mse_use_sigmoid=True, **training_defaults):
        super().__init__()
        assert recon_loss in {"bce", "mse"}, f"recon_loss must be 'bce' or 'mse', got {recon_loss}"
        self.encoder = Encoder(in_channels, enc_channels, latent_dim, norm_type)
        self.decoder = Decoder(in_channels, dec_channels, latent_dim, output_image_size, norm_type)
        self.recon_loss, self.mse_use_sigmoid = recon_loss, mse_use_sigmoid
        self.training_defaults = dict(training_defaults)  # free-bits / capacity knobs of the (out-of-scope) trainer

    def forward(self, x):
        mu, logvar = self.encoder(x)
        z = mu + torch.randn_like(mu) * torch.exp(0.5 * logvar)
        return self.decoder(z), mu, logvar, z
