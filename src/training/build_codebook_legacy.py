"""Same path as the reference's src/training/build_codebook_legacy.py; forwards to vqvae_b200."""
from vqvae_b200.training.build_codebook_legacy import *  # noqa: F401,F403
from vqvae_b200.training.build_codebook_legacy import build_and_save  # noqa: F401

if __name__ == "__main__":
    import runpy
    runpy.run_module("vqvae_b200.training.build_codebook_legacy", run_name="__main__")
