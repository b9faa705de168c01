from vqvae_b200.geo.geo_shortest_paths import *  # noqa: F401,F403
from vqvae_b200.geo import geo_shortest_paths as _impl

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith("__")})
