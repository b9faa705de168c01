"""Same path as the reference's CLI (src/scripts/build_codebook.py); forwards to vqvae_b200."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from vqvae_b200.scripts.build_codebook import main, make_parser  # noqa: E402

if __name__ == "__main__":
    main(make_parser().parse_args())
