from vqvae_b200.geo.kmeans_optimized import *  # noqa: F401,F403
from vqvae_b200.geo import kmeans_optimized as _impl

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith("__")})
