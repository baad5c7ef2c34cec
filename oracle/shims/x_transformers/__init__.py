"""Shim of the third-party `x_transformers` package (>=1.31.14, not installed here).  TEST INFRASTRUCTURE.
Only the two RoPE symbols the reference imports are restated, from the package's published algorithm
(call sites: reference backbones/dit.py:17,245,454; model/modules.py:20,470-480)."""
from .x_transformers import RotaryEmbedding, apply_rotary_pos_emb, RMSNorm  # noqa: F401
