from vqvae_b200.geo.riemannian_metric import *  # noqa: F401,F403
from vqvae_b200.geo import riemannian_metric as _impl

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith("__")})
