import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "oracle", ROOT / "tests" / "golden"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

FASHION_CFG = dict(in_channels=1, output_image_size=28, latent_dim=16, enc_channels=[64, 128, 256],
                   dec_channels=[256, 128, 64], recon_loss="mse", norm_type="batch", mse_use_sigmoid=True)


CIFAR_CFG = dict(in_channels=3, output_image_size=32, latent_dim=32, enc_channels=[64, 128, 256],
                 dec_channels=[256, 128, 64], recon_loss="mse", norm_type="batch", mse_use_sigmoid=False)
for p in (ROOT / "tools",):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return ROOT / "tests" / "golden"


@pytest.fixture(scope="session")
def s1(golden_dir):
    import numpy as np
    return dict(np.load(golden_dir / "s1_stages.npz"))


@pytest.fixture(scope="session")
def c1(golden_dir):
    import numpy as np
    return dict(np.load(golden_dir / "c1_digest.npz"))


@pytest.fixture(scope="session")
def c3(golden_dir):
    import numpy as np
    return dict(np.load(golden_dir / "c3_digest.npz"))


def c3_latents():
    import torch
    z4 = torch.randn(1250, 32, 4, 4, generator=torch.Generator().manual_seed(42))
    return z4, z4.permute(0, 2, 3, 1).reshape(-1, 32).numpy()


def cifar_decoder():
    import torch
    from vqvae_b200.models import SpatialVAE
    torch.manual_seed(0)
    return SpatialVAE(**CIFAR_CFG).decoder  # train mode, like the reference CLI


def s1_latents():
    import torch
    z4 = torch.randn(125, 16, 4, 4, generator=torch.Generator().manual_seed(42))
    return z4, z4.permute(0, 2, 3, 1).reshape(-1, 16).numpy()


def c1_latents():
    import torch
    z4 = torch.randn(1250, 16, 4, 4, generator=torch.Generator().manual_seed(42))
    return z4, z4.permute(0, 2, 3, 1).reshape(-1, 16).numpy()


def fashion_decoder():
    import torch
    from vqvae_b200.models import SpatialVAE
    torch.manual_seed(0)
    return SpatialVAE(**FASHION_CFG).decoder  # train mode, like the reference CLI


def csr_from(indptr, indices, data, n):
    from scipy import sparse
    return sparse.csr_matrix((data, indices, indptr), shape=(n, n))
