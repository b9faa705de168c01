from vqvae_b200.geo.knn_graph_optimized import *  # noqa: F401,F403
from vqvae_b200.geo import knn_graph_optimized as _impl

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith("__")})
