"""Import shim: lets reference-side code (`from src.geo... import ...`,
`python src/scripts/build_codebook.py`) run against vqvae_b200 unchanged."""
