"""CPU/torch restatement of the F5E-TTS inference hot path.  TEST INFRASTRUCTURE ONLY.

This file is the ORACLE: a plain, unfused torch restatement (any device / dtype, fp32 by default) of the
reference algorithm, written from the reference's behaviour with each function citing the reference
file:line it follows (paths relative to /root/reference/src/f5_tts/).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import it; the product (f5e_tts_b200/) never does.

Pinning: the reference ships NO tests or golden vectors for this path, so the oracle is
pinned against outputs of the reference itself, imported in the build container through oracle/shims/
(oracle/make_golden.py -> tests/golden/*.pt, checked by tests/test_oracle_golden.py).  The Vocos backbone
and the torchdiffeq / x_transformers maths live in third-party packages that are absent from the reference
tree (pyproject.toml:37,41,43, unpinned): those are restated from their published algorithms and that part
of the parity is "unpinned" (only the ISTFT head is corroborated in-tree:
runtime/triton_trtllm/scripts/export_vocoder_to_onnx.py:45-59).

State-dict keys are exactly the reference's (`transformer.` prefix stripped), so a reference checkpoint
loads into the oracle and into the product unchanged.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch
import torch.nn.functional as F


@dataclass
class DiTConfig:
    """Constructor arguments of reference backbones/dit.py:183-204 that matter at inference."""
    dim: int = 1024
    depth: int = 22
    heads: int = 16
    dim_head: int = 64
    ff_mult: int = 2
    mel_dim: int = 100
    text_num_embeds: int = 2545
    text_dim: int = 512
    text_mask_padding: bool = True
    conv_layers: int = 4
    pe_attn_head: int | None = None
    use_ppg: bool = False
    ppg_dim: int = 256
    use_codebook: bool = False
    cb_num_vars: int = 100
    cb_groups: int = 2

    @property
    def in_dim(self) -> int:
        return self.mel_dim * 2 + self.text_dim * (2 if self.use_ppg else 1)


F5TTS_V1_BASE = DiTConfig()
# configs/example.yaml:25-34, 59-67 (the PPG + codebook "Small" F5E variant, BASELINE config 3)
F5E_SMALL_PPG = DiTConfig(dim=768, depth=18, heads=12, text_mask_padding=False, pe_attn_head=1, use_ppg=True,
                          use_codebook=True)


# -------------------------------------------------------------------------------------------------------
# weights
# -------------------------------------------------------------------------------------------------------
def make_state_dict(cfg: DiTConfig, seed: int = 1234, device="cpu", zero_init_std: float = 0.02) -> dict:
    """Random weights in the reference's state-dict layout.  Linear/Conv use torch's default
    uniform(-1/sqrt(fan_in), 1/sqrt(fan_in)); the layers the reference zero-initialises
    (dit.py:273-283: adaLN linears, norm_out.linear, proj_out) get N(0, zero_init_std) instead, because
    with zeros the model output is identically zero and parity would be vacuous."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd: dict[str, torch.Tensor] = {}

    def uni(name, shape, fan_in):
        b = 1.0 / math.sqrt(fan_in)
        sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * b

    def lin(prefix, out_f, in_f, std=None):
        if std is None:
            uni(prefix + ".weight", (out_f, in_f), in_f)
            uni(prefix + ".bias", (out_f,), in_f)
        else:
            sd[prefix + ".weight"] = torch.randn(out_f, in_f, generator=g) * std
            sd[prefix + ".bias"] = torch.randn(out_f, generator=g) * std

    D, TD = cfg.dim, cfg.text_dim
    lin("time_embed.time_mlp.0", D, 256)
    lin("time_embed.time_mlp.2", D, D)
    sd["text_embed.text_embed.weight"] = torch.randn(cfg.text_num_embeds + 1, TD, generator=g)
    for i in range(cfg.conv_layers):
        p = f"text_embed.text_blocks.{i}"
        uni(p + ".dwconv.weight", (TD, 1, 7), 7)
        uni(p + ".dwconv.bias", (TD,), 7)
        sd[p + ".norm.weight"] = 1 + 0.1 * torch.randn(TD, generator=g)
        sd[p + ".norm.bias"] = 0.1 * torch.randn(TD, generator=g)
        lin(p + ".pwconv1", TD * 2, TD)
        sd[p + ".grn.gamma"] = 0.1 * torch.randn(1, 1, TD * 2, generator=g)  # reference inits zeros
        sd[p + ".grn.beta"] = 0.1 * torch.randn(1, 1, TD * 2, generator=g)
        lin(p + ".pwconv2", TD, TD * 2)
    if cfg.use_ppg:
        P = cfg.ppg_dim
        lin("ppg_embed.ppg_proj.0", P, P)
        for ci, bi in ((2, 3), (6, 7), (10, 11)):
            uni(f"ppg_embed.ppg_proj.{ci}.weight", (P, P, 5), P * 5)
            uni(f"ppg_embed.ppg_proj.{ci}.bias", (P,), P * 5)
            sd[f"ppg_embed.ppg_proj.{bi}.weight"] = 1 + 0.1 * torch.randn(P, generator=g)
            sd[f"ppg_embed.ppg_proj.{bi}.bias"] = 0.1 * torch.randn(P, generator=g)
            sd[f"ppg_embed.ppg_proj.{bi}.running_mean"] = 0.1 * torch.randn(P, generator=g)
            sd[f"ppg_embed.ppg_proj.{bi}.running_var"] = 1 + 0.2 * torch.rand(P, generator=g)
        lin("ppg_embed.ppg_proj.15", TD, P)
    if cfg.use_codebook:
        sd["quantizer.vars"] = torch.rand(1, cfg.cb_groups * cfg.cb_num_vars, TD // cfg.cb_groups, generator=g)
        sd["quantizer.weight_proj.weight"] = torch.randn(cfg.cb_groups * cfg.cb_num_vars, TD, generator=g)
        sd["quantizer.weight_proj.bias"] = torch.zeros(cfg.cb_groups * cfg.cb_num_vars)
    lin("input_embed.proj", D, cfg.in_dim)
    cpg = D // 16
    for i in (0, 2):
        uni(f"input_embed.conv_pos_embed.conv1d.{i}.weight", (D, cpg, 31), cpg * 31)
        uni(f"input_embed.conv_pos_embed.conv1d.{i}.bias", (D,), cpg * 31)
    inner = cfg.heads * cfg.dim_head
    for l in range(cfg.depth):
        p = f"transformer_blocks.{l}"
        lin(p + ".attn_norm.linear", 6 * D, D, std=zero_init_std)
        lin(p + ".attn.to_q", inner, D)
        lin(p + ".attn.to_k", inner, D)
        lin(p + ".attn.to_v", inner, D)
        lin(p + ".attn.to_out.0", D, inner)
        lin(p + ".ff.ff.0.0", D * cfg.ff_mult, D)
        lin(p + ".ff.ff.2", D, D * cfg.ff_mult)
    lin("norm_out.linear", 2 * D, D, std=zero_init_std)
    lin("proj_out", cfg.mel_dim, D, std=zero_init_std)
    return {k: v.to(device) for k, v in sd.items()}
