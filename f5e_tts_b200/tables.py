"""Constant tables of the path (position sinusoids, RoPE angles, mel filterbank, windows).  Built once per plan with
torch on the target device; they are inputs of the kernels, not compute on the hot path."""
from __future__ import annotations

import math

import torch


def text_pos_table(dim: int, end: int = 4096, theta: float = 10000.0) -> torch.Tensor:
    """fp32 [end, dim] = [cos(p f_j) | sin(p f_j)]  (reference model/modules.py:196-207)."""
    freqs = 1.0 / (theta ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
    f = torch.outer(torch.arange(end), freqs).float()
    return torch.cat([f.cos(), f.sin()], dim=-1).contiguous()
