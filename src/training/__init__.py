"""`src.training`: the two legacy YAML codebook drivers are this repo's (vqvae_b200.training); the reference's other
training modules (VAE / Transformer engines, out of scope) resolve in its checkout when one is reachable."""
from .. import overlay as _overlay

_overlay(__path__, "training")
