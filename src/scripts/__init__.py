"""`src.scripts`: build_codebook is this repo's CLI; the reference's other scripts resolve in its checkout."""
from .. import overlay as _overlay

_overlay(__path__, "scripts")
