"""Downstream consumers of the codebook artifacts that touch the hot path."""
from .nearest_medoid import nearest_medoid_assign, decode_codes, valid_code_rows  # noqa: F401
