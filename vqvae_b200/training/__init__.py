"""Legacy YAML-driven codebook builders of the reference (flat "vanilla" latents; BASELINE configs[3])."""
