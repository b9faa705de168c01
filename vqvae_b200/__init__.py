"""vqvae_b200 -- B200-native geodesic codebook builder (drop-in for m4rch1n0/vqvae's
`src/geo` + `src/scripts/build_codebook.py` path).  Hot ops are sm_100a CUDA kernels in
`csrc/` behind the C ABI of `include/geocb.h`; no CPU fallback exists."""
__version__ = "0.1.0"
