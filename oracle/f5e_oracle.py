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
    qk_norm: str | None = None          # "rms_norm": RMSNorm(dim_head) on q and k before RoPE (modules.py:391-398, 463-467)
    long_skip_connection: bool = False  # Linear(2 dim -> dim, no bias) over cat(x, x_before_the_blocks) (dit.py:264, 456-467)
    use_ppg: bool = False
    ppg_dim: int = 256
    use_codebook: bool = False
    cb_num_vars: int = 100
    cb_groups: int = 2
    # PPGEmbedding(use_transformer=True) (dit.py:105-119): nn.TransformerEncoder over ppg_dim, then Linear -> text_dim
    ppg_transformer: bool = False
    ppg_nhead: int = 4
    ppg_ff: int = 1024
    ppg_layers: int = 2

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
    if cfg.use_ppg and cfg.ppg_transformer:
        P = cfg.ppg_dim
        for i in range(cfg.ppg_layers):
            p = f"ppg_embed.ppg_proj.0.layers.{i}"
            uni(p + ".self_attn.in_proj_weight", (3 * P, P), P)
            uni(p + ".self_attn.in_proj_bias", (3 * P,), P)
            lin(p + ".self_attn.out_proj", P, P)
            lin(p + ".linear1", cfg.ppg_ff, P)
            lin(p + ".linear2", P, cfg.ppg_ff)
            for nm in ("norm1", "norm2"):
                sd[p + f".{nm}.weight"] = 1 + 0.1 * torch.randn(P, generator=g)
                sd[p + f".{nm}.bias"] = 0.1 * torch.randn(P, generator=g)
        lin("ppg_embed.ppg_proj.1", TD, P)
    elif cfg.use_ppg:
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
        if cfg.qk_norm == "rms_norm":  # the reference initialises these to ones; random so that a wrong weight shows
            sd[p + ".attn.q_norm.weight"] = 1.0 + 0.3 * torch.randn(cfg.dim_head, generator=g)
            sd[p + ".attn.k_norm.weight"] = 1.0 + 0.3 * torch.randn(cfg.dim_head, generator=g)
    if cfg.long_skip_connection:
        sd["long_skip_connection.weight"] = torch.randn(D, 2 * D, generator=g) / (2 * D) ** 0.5
    lin("norm_out.linear", 2 * D, D, std=zero_init_std)
    lin("proj_out", cfg.mel_dim, D, std=zero_init_std)
    return {k: v.to(device) for k, v in sd.items()}


# -------------------------------------------------------------------------------------------------------
# front end: MelSpec  (model/modules.py:75-143 == torchaudio MelSpectrogram, restated with torch.stft;
# max-abs diff 0.0 against torchaudio)
# -------------------------------------------------------------------------------------------------------
def hz_to_mel_htk(f):
    return 2595.0 * math.log10(1.0 + f / 700.0)


def mel_filterbank(n_freqs=513, f_min=0.0, f_max=12000.0, n_mels=100, sample_rate=24000) -> torch.Tensor:
    """HTK mel filterbank, norm=None, [n_freqs, n_mels] (torchaudio.functional.melscale_fbanks algorithm)."""
    all_freqs = torch.linspace(0, sample_rate // 2, n_freqs)
    m_pts = torch.linspace(hz_to_mel_htk(f_min), hz_to_mel_htk(f_max), n_mels + 2)
    f_pts = 700.0 * (10.0 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    down = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    return torch.clamp(torch.min(down, up), min=0.0)


def mel_spectrogram(wav: torch.Tensor, n_fft=1024, hop=256, win=1024, n_mels=100, sr=24000) -> torch.Tensor:
    """wav [B, nw] fp32 -> log-mel [B, n_mels, nw // hop + 1]."""
    if wav.dim() == 3:
        wav = wav.squeeze(1)
    window = torch.hann_window(win, periodic=True, device=wav.device, dtype=wav.dtype)
    spec = torch.stft(wav, n_fft, hop_length=hop, win_length=win, window=window, center=True, pad_mode="reflect",
                      normalized=False, onesided=True, return_complex=True).abs()
    fb = mel_filterbank(n_fft // 2 + 1, 0.0, sr / 2, n_mels, sr).to(wav.device, wav.dtype)
    mel = torch.matmul(spec.transpose(-1, -2), fb).transpose(-1, -2)
    return mel.clamp(min=1e-5).log()


# -------------------------------------------------------------------------------------------------------
# DiT evaluation  (backbones/dit.py:417-472)
# -------------------------------------------------------------------------------------------------------
def _lin(sd, p, x):
    return F.linear(x, sd[p + ".weight"], sd[p + ".bias"])


def sinus_time_embed(t: torch.Tensor, dim=256, scale=1000) -> torch.Tensor:
    """model/modules.py:149-161, incl. its dtype behaviour: `scale * t` happens in t's dtype, the rest in fp32."""
    half = dim // 2
    e = math.log(10000) / (half - 1)
    e = torch.exp(torch.arange(half, device=t.device).float() * -e)
    e = scale * t.unsqueeze(1) * e.unsqueeze(0)
    return torch.cat((e.sin(), e.cos()), dim=-1)


def text_pos_table(dim: int, end: int = 4096, theta: float = 10000.0) -> torch.Tensor:
    """model/modules.py:196-207 -> [end, dim] = [cos | sin]."""
    freqs = 1.0 / (theta ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
    f = torch.outer(torch.arange(end), freqs).float()
    return torch.cat([f.cos(), f.sin()], dim=-1)


def rope_angles(n: int, dim_head=64, device="cpu") -> torch.Tensor:
    """x_transformers RotaryEmbedding.forward_from_seq_len: interleaved pairs, [n, dim_head] fp32."""
    inv = 1.0 / (10000 ** (torch.arange(0, dim_head, 2, device=device).float() / dim_head))
    fr = torch.arange(n, device=device).float()[:, None] * inv[None]
    return torch.stack((fr, fr), dim=-1).reshape(n, dim_head)


def apply_rope(t: torch.Tensor, ang: torch.Tensor) -> torch.Tensor:
    """x_transformers apply_rotary_pos_emb: fp32 math, cast back (modules.py:470-480 call site)."""
    dt = t.dtype
    x = t.float().reshape(*t.shape[:-1], -1, 2)
    rot = torch.stack((-x[..., 1], x[..., 0]), dim=-1).reshape(t.shape)
    return (t.float() * ang.cos() + rot * ang.sin()).to(dt)


class OracleDiT:
    """Functional restatement of DiT.sample over a reference-layout state dict."""

    def __init__(self, cfg: DiTConfig, sd: dict, dtype=torch.float32, device="cpu"):
        self.cfg = cfg
        self.dtype = dtype
        self.device = torch.device(device)
        self.sd = {k: (v.to(self.device, dtype) if v.is_floating_point() else v.to(self.device)) for k, v in sd.items()}
        # non-persistent buffer, cast with the module (model.to(dtype) casts float buffers)
        self.pos = text_pos_table(cfg.text_dim).to(self.device, dtype)
        self.text_cond = self.text_uncond = None

    def clear_cache(self):
        self.text_cond = self.text_uncond = None

    # -- TextEmbedding (dit.py:54-87) + ConvNeXtV2Block (modules.py:241-269) + GRN (:225-234)
    def text_embed(self, text, batch, seq_len, drop_text):
        sd, cfg = self.sd, self.cfg
        if text is None:  # dit.py:55-56 (the VC path); the reference has no text_mask here, so mask_padding must be off
            assert not cfg.text_mask_padding, "reference raises NameError: text=None needs text_mask_padding=False"
            text = torch.zeros((batch, seq_len), dtype=torch.long, device=self.device)
            text_mask = None
        else:
            text = text + 1
            text = text[:, :seq_len]
            text = F.pad(text, (0, seq_len - text.shape[1]), value=0)
            text_mask = text == 0
            if drop_text:
                text = torch.zeros_like(text)
        x = F.embedding(text, sd["text_embed.text_embed.weight"])
        if cfg.conv_layers > 0:
            pos_idx = torch.arange(seq_len, device=x.device).clamp(max=4095)
            x = x + self.pos[pos_idx][None]
            mask = text_mask.unsqueeze(-1) if cfg.text_mask_padding else None
            if mask is not None:
                x = x.masked_fill(mask, 0.0)
            for i in range(cfg.conv_layers):
                p = f"text_embed.text_blocks.{i}"
                r = x
                h = F.conv1d(x.transpose(1, 2), sd[p + ".dwconv.weight"], sd[p + ".dwconv.bias"], padding=3,
                             groups=cfg.text_dim).transpose(1, 2)
                h = F.layer_norm(h, (cfg.text_dim,), sd[p + ".norm.weight"], sd[p + ".norm.bias"], eps=1e-6)
                h = F.gelu(_lin(sd, p + ".pwconv1", h))
                gx = torch.norm(h, p=2, dim=1, keepdim=True)
                nx = gx / (gx.mean(dim=-1, keepdim=True) + 1e-6)
                h = sd[p + ".grn.gamma"] * (h * nx) + sd[p + ".grn.beta"] + h
                x = r + _lin(sd, p + ".pwconv2", h)
                if mask is not None:
                    x = x.masked_fill(mask, 0.0)
        return x

    # -- PPGEmbedding (dit.py:121-153), eval-mode BatchNorm
    def ppg_embed(self, ppg, seq_len, drop_ppg, batch):
        sd, cfg = self.sd, self.cfg
        if ppg is None:
            ppg = torch.zeros(batch, seq_len, cfg.ppg_dim, dtype=self.dtype, device=self.device)
        else:
            ppg = F.pad(ppg, (0, 0, 0, seq_len - ppg.shape[1]), value=0).to(self.dtype)
            if drop_ppg:
                ppg = torch.zeros_like(ppg)
        if cfg.ppg_transformer:  # nn.TransformerEncoderLayer, post-norm, exact GELU, eval mode, no padding mask (dit.py:105-119)
            x, P, H = ppg, cfg.ppg_dim, cfg.ppg_nhead
            B, N = x.shape[:2]
            for i in range(cfg.ppg_layers):
                p = f"ppg_embed.ppg_proj.0.layers.{i}"
                qkv = F.linear(x, sd[p + ".self_attn.in_proj_weight"], sd[p + ".self_attn.in_proj_bias"])
                q, k, v = (t.reshape(B, N, H, P // H).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
                a = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, N, P)
                x = F.layer_norm(x + _lin(sd, p + ".self_attn.out_proj", a), (P,), sd[p + ".norm1.weight"], sd[p + ".norm1.bias"], 1e-5)
                f = _lin(sd, p + ".linear2", F.gelu(_lin(sd, p + ".linear1", x)))
                x = F.layer_norm(x + f, (P,), sd[p + ".norm2.weight"], sd[p + ".norm2.bias"], 1e-5)
            return _lin(sd, "ppg_embed.ppg_proj.1", x)
        h = _lin(sd, "ppg_embed.ppg_proj.0", ppg).transpose(1, 2)
        for ci, bi in ((2, 3), (6, 7), (10, 11)):
            h = F.conv1d(h, sd[f"ppg_embed.ppg_proj.{ci}.weight"], sd[f"ppg_embed.ppg_proj.{ci}.bias"], padding=2)
            h = F.batch_norm(h, sd[f"ppg_embed.ppg_proj.{bi}.running_mean"], sd[f"ppg_embed.ppg_proj.{bi}.running_var"],
                             sd[f"ppg_embed.ppg_proj.{bi}.weight"], sd[f"ppg_embed.ppg_proj.{bi}.bias"], False, 0.0, 1e-5)
            h = F.relu(h)
        return _lin(sd, "ppg_embed.ppg_proj.15", h.transpose(1, 2))

    def sample(self, x, cond, text, ppg, time, drop_audio_cond, drop_text, drop_ppg, mask=None, taps=None):
        sd, cfg = self.sd, self.cfg
        B, N = x.shape[0], x.shape[1]
        if time.ndim == 0:
            time = time.repeat(B)
        # TimestepEmbedding (modules.py:721-731)
        th = sinus_time_embed(time).to(time.dtype)
        t = _lin(sd, "time_embed.time_mlp.2", F.silu(_lin(sd, "time_embed.time_mlp.0", th)))
        if drop_text:
            if self.text_uncond is None:
                self.text_uncond = self.text_embed(text, B, N, True)
            te = self.text_uncond
        else:
            if self.text_cond is None:
                self.text_cond = self.text_embed(text, B, N, False)
            te = self.text_cond
        if drop_audio_cond:
            cond = torch.zeros_like(cond)
        parts = [x, cond, te]
        if cfg.use_ppg:
            parts.append(self.ppg_embed(ppg, N, drop_ppg, B))
        h = _lin(sd, "input_embed.proj", torch.cat(parts, dim=-1))
        # ConvPositionEmbedding, unmasked (modules.py:167-190; dit.py:176)
        c = h.transpose(1, 2)
        for i in (0, 2):
            c = F.mish(F.conv1d(c, sd[f"input_embed.conv_pos_embed.conv1d.{i}.weight"],
                                sd[f"input_embed.conv_pos_embed.conv1d.{i}.bias"], padding=15, groups=16))
        h = c.transpose(1, 2) + h
        if taps is not None:
            taps["embed"] = h
        ang = rope_angles(N, cfg.dim_head, x.device)
        H, dh = cfg.heads, cfg.dim_head
        ts = F.silu(t)
        residual = h  # dit.py:456-457 (used only with long_skip_connection)
        for l in range(cfg.depth):
            p = f"transformer_blocks.{l}"
            emb = _lin(sd, p + ".attn_norm.linear", ts)
            shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp = torch.chunk(emb, 6, dim=1)
            nrm = F.layer_norm(h, (cfg.dim,), eps=1e-6) * (1 + scale_msa[:, None]) + shift_msa[:, None]
            q = _lin(sd, p + ".attn.to_q", nrm).view(B, N, H, dh).transpose(1, 2)
            k = _lin(sd, p + ".attn.to_k", nrm).view(B, N, H, dh).transpose(1, 2)
            v = _lin(sd, p + ".attn.to_v", nrm).view(B, N, H, dh).transpose(1, 2)
            if cfg.qk_norm == "rms_norm":  # modules.py:463-467, RMSNorm.forward :282-294 (F.rms_norm over dim_head)
                q = F.rms_norm(q, (dh,), weight=sd[p + ".attn.q_norm.weight"], eps=1e-6)
                k = F.rms_norm(k, (dh,), weight=sd[p + ".attn.k_norm.weight"], eps=1e-6)
            pn = H if cfg.pe_attn_head is None else cfg.pe_attn_head
            q = torch.cat((apply_rope(q[:, :pn], ang), q[:, pn:]), dim=1)
            k = torch.cat((apply_rope(k[:, :pn], ang), k[:, pn:]), dim=1)
            if mask is not None and getattr(self, "attn_backend", "mask") == "flash":
                # timing arm only (bench.py gpu_baseline): what an eager implementation free to pick the flash SDPA backend
                # would run - per sample over its valid frames; padded query rows are zeroed by the masked_fill below anyway
                a = torch.zeros_like(q)
                for i, L in enumerate(mask.sum(-1).tolist()):
                    a[i, :, :L] = F.scaled_dot_product_attention(q[i:i + 1, :, :L], k[i:i + 1, :, :L], v[i:i + 1, :, :L])[0]
            else:
                am = None if mask is None else mask[:, None, None, :].expand(B, H, N, N)
                a = F.scaled_dot_product_attention(q, k, v, attn_mask=am)
            a = a.transpose(1, 2).reshape(B, N, H * dh)
            a = _lin(sd, p + ".attn.to_out.0", a)
            if mask is not None:
                a = a.masked_fill(~mask.unsqueeze(-1), 0.0)
            h = h + gate_msa.unsqueeze(1) * a
            nrm = F.layer_norm(h, (cfg.dim,), eps=1e-6) * (1 + scale_mlp[:, None]) + shift_mlp[:, None]
            ff = _lin(sd, p + ".ff.ff.2", F.gelu(_lin(sd, p + ".ff.ff.0.0", nrm), approximate="tanh"))
            h = h + gate_mlp.unsqueeze(1) * ff
            if taps is not None and l == 0:
                taps["block0"] = h
        if cfg.long_skip_connection:  # dit.py:466-467
            h = F.linear(torch.cat((h, residual), dim=-1), sd["long_skip_connection.weight"])
        emb = _lin(sd, "norm_out.linear", ts)
        scale, shift = torch.chunk(emb, 2, dim=1)
        h = F.layer_norm(h, (cfg.dim,), eps=1e-6) * (1 + scale)[:, None, :] + shift[:, None, :]
        return _lin(sd, "proj_out", h)


# -------------------------------------------------------------------------------------------------------
# CFM.sample  (model/cfm.py:348-482) with torchdiffeq's fixed-grid Euler restated inline
# -------------------------------------------------------------------------------------------------------
def lens_to_mask(t: torch.Tensor, length=None) -> torch.Tensor:
    """model/utils.py:41-46."""
    if length is None:
        length = int(t.amax())
    return torch.arange(length, device=t.device)[None, :] < t[:, None]


def sway_schedule(steps, sway, dtype, device, t_start=0.0):
    """model/cfm.py:467-469, in the model dtype like the reference."""
    t = torch.linspace(t_start, 1, steps + 1, device=device, dtype=dtype)
    if sway is not None:
        t = t + sway * (torch.cos(torch.pi / 2 * t) - 1 + t)
    return t


@torch.no_grad()
def cfm_sample(dit: OracleDiT, cond, text, duration, *, ppg=None, lens=None, steps=32, cfg_strength=1.0,
               sway_sampling_coef=None, seed=None, max_duration=4096, no_ref_audio=False, edit_mask=None,
               duplicate_test=False, t_inter=0.1, vocoder=None,
               y0=None, return_aux=False, mode="cfg", alpha_spk=1.0, alpha_cont=1.0):
    """cond: mel [B, T, 100] (or wave [B, nw]); text: int64 [B, nt] padded with -1.
    Returns (out, trajectory) like the reference; with return_aux also the integer/bool side products.
    mode "cfg" = CFM.sample (model/cfm.py:348-482); "tts" = CFM.sample_tts (:93-223, text content, ppg must be None,
    alpha_cont = alpha_txt); "vc" = CFM.sample_vc (:225-346, PPG content, text must be None, alpha_cont = alpha_ppg)."""
    assert mode in ("cfg", "tts", "vc")
    if mode == "tts":
        assert ppg is None
    if mode == "vc":
        assert text is None
    dtype, device = dit.dtype, dit.device
    if cond.ndim == 2:
        cond = mel_spectrogram(cond).permute(0, 2, 1)
    cond = cond.to(dtype)
    batch, cond_seq_len = cond.shape[:2]
    if lens is None:
        lens = torch.full((batch,), cond_seq_len, device=device, dtype=torch.long)
    cond_mask = lens_to_mask(lens)
    if edit_mask is not None:
        cond_mask = cond_mask & edit_mask
    if isinstance(duration, int):
        duration = torch.full((batch,), duration, device=device, dtype=torch.long)
    if text is not None:
        duration = torch.maximum(torch.maximum((text != -1).sum(dim=-1), lens) + 1, duration)
    else:
        duration = torch.maximum(lens + 1, duration)
    duration = duration.clamp(max=max_duration)
    max_dur = int(duration.amax())
    if duplicate_test:  # cfm.py:410-412: the prompt repeated right after itself, blended into the noise below
        test_cond = F.pad(cond, (0, 0, cond_seq_len, max_dur - 2 * cond_seq_len), value=0.0)
    cond = F.pad(cond, (0, 0, 0, max_dur - cond_seq_len), value=0.0)
    if no_ref_audio:
        cond = torch.zeros_like(cond)
    cond_mask = F.pad(cond_mask, (0, max_dur - cond_mask.shape[-1]), value=False).unsqueeze(-1)
    step_cond = torch.where(cond_mask, cond, torch.zeros_like(cond))
    mask = lens_to_mask(duration) if batch > 1 else None

    def fn(t, x):
        if mode != "cfg":  # three evaluations: null, content only, speaker + content (cfm.py:166-185, :298-317)
            keep_txt, keep_ppg = mode == "tts", mode == "vc"
            null_flow = dit.sample(x, step_cond, text, ppg, t, True, True, True, mask)
            cont_flow = dit.sample(x, step_cond, text, ppg, t, True, not keep_txt, not keep_ppg, mask)
            full_flow = dit.sample(x, step_cond, text, ppg, t, False, not keep_txt, not keep_ppg, mask)
            return alpha_spk * (full_flow - cont_flow) + alpha_cont * (cont_flow - null_flow) + null_flow
        pred = dit.sample(x, step_cond, text, ppg, t, False, False, False, mask)
        if cfg_strength < 1e-5:
            return pred
        null_pred = dit.sample(x, step_cond, text, ppg, t, True, True, True, mask)
        return pred + (pred - null_pred) * cfg_strength

    if y0 is None:
        ys = []
        for dur in duration:
            if seed is not None:
                torch.manual_seed(seed)
            ys.append(torch.randn(int(dur), cond.shape[-1], device=device, dtype=dtype))
        y0 = torch.nn.utils.rnn.pad_sequence(ys, padding_value=0, batch_first=True)
    t_start = 0
    if duplicate_test:  # cfm.py:460-465
        t_start = t_inter
        y0 = (1 - t_start) * y0 + t_start * test_cond
        steps = int(steps * (1 - t_start))
    t = sway_schedule(steps, sway_sampling_coef, dtype, device, t_start)
    traj = [y0]
    y = y0
    for i in range(steps):  # torchdiffeq fixed-grid Euler: y1 = y0 + (t1 - t0) * f(t0, y0)
        dt = t[i + 1] - t[i]
        y = y + dt * fn(t[i], y)
        traj.append(y)
    dit.clear_cache()
    trajectory = torch.stack(traj)
    out = torch.where(cond_mask, cond, trajectory[-1])
    if vocoder is not None:  # cfm.py:478-480
        out = vocoder(out.permute(0, 2, 1))
    if return_aux:
        return out, trajectory, dict(duration=duration, cond_mask=cond_mask, mask=mask, t=t, y0=y0, step_cond=step_cond)
    return out, trajectory


# -------------------------------------------------------------------------------------------------------
# GumbelVectorQuantizer.forward in eval mode  (model/modules.py:881-950)
# -------------------------------------------------------------------------------------------------------
def vq_forward(sd: dict, x: torch.Tensor, groups=2, num_vars=100) -> dict:
    bsz, tsz, fsz = x.shape
    w, b, vars_ = sd["quantizer.weight_proj.weight"], sd["quantizer.weight_proj.bias"], sd["quantizer.vars"]
    logits = F.linear(x.reshape(-1, fsz), w, b).view(bsz * tsz * groups, -1)
    _, k = logits.max(-1)
    hard_x = logits.new_zeros(*logits.shape).scatter_(-1, k.view(-1, 1), 1.0).view(bsz * tsz, groups, -1)
    hard_probs = torch.mean(hard_x.float(), dim=0)
    code_ppl = torch.exp(-torch.sum(hard_probs * torch.log(hard_probs + 1e-7), dim=-1)).sum()
    avg_probs = torch.softmax(logits.view(bsz * tsz, groups, -1).float(), dim=-1).mean(dim=0)
    prob_ppl = torch.exp(-torch.sum(avg_probs * torch.log(avg_probs + 1e-7), dim=-1)).sum()
    targets = hard_x.view(bsz * tsz * groups, -1).argmax(dim=-1).view(bsz, tsz, groups)
    xq = hard_x.view(bsz * tsz, -1).unsqueeze(-1) * vars_
    xq = xq.view(bsz * tsz, groups, num_vars, -1).sum(-2).view(bsz, tsz, -1)
    return {"x": xq, "targets": targets, "num_vars": num_vars * groups, "code_perplexity": code_ppl,
            "prob_perplexity": prob_ppl}


# -------------------------------------------------------------------------------------------------------
# Vocos mel-24khz decode.  PARITY UNPINNED for the backbone: the `vocos` package is an unpinned third-party dependency
# (pyproject.toml:41) that is neither vendored nor installed, and no reference test touches it; the backbone below is
# restated from the package's published architecture (VocosBackbone + ConvNeXtBlock + ISTFTHead, config of
# charactr/vocos-mel-24khz).  The head is corroborated in-tree: runtime/triton_trtllm/scripts/export_vocoder_to_onnx.py:45-59
# (exp / clip 1e2 / cos / sin) and scripts/conv_stft.py:193-234 (window, overlap-add, envelope); SURVEY appendix A pins
# ISTFT(padding="center") == torch.istft(center=True) numerically.
# -------------------------------------------------------------------------------------------------------
def make_vocos_state_dict(seed: int = 4321, in_ch=100, dim=512, inter=1536, n_layers=8, n_fft=1024) -> dict:
    """Random weights in the package's state-dict layout with its init (trunc_normal std 0.02, zero bias,
    gamma = 1 / n_layers); norm affine parameters are perturbed so that they are exercised."""
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def tn(*shape):
        return torch.nn.init.trunc_normal_(torch.empty(*shape), std=0.02, generator=g)

    sd["backbone.embed.weight"], sd["backbone.embed.bias"] = tn(dim, in_ch, 7), 0.02 * torch.randn(dim, generator=g)
    sd["backbone.norm.weight"], sd["backbone.norm.bias"] = 1 + 0.1 * torch.randn(dim, generator=g), 0.1 * torch.randn(dim, generator=g)
    for i in range(n_layers):
        p = f"backbone.convnext.{i}"
        sd[p + ".dwconv.weight"], sd[p + ".dwconv.bias"] = tn(dim, 1, 7) * 10, 0.02 * torch.randn(dim, generator=g)
        sd[p + ".norm.weight"], sd[p + ".norm.bias"] = 1 + 0.1 * torch.randn(dim, generator=g), 0.1 * torch.randn(dim, generator=g)
        sd[p + ".pwconv1.weight"], sd[p + ".pwconv1.bias"] = tn(inter, dim) * 3, 0.02 * torch.randn(inter, generator=g)
        sd[p + ".pwconv2.weight"], sd[p + ".pwconv2.bias"] = tn(dim, inter) * 3, 0.02 * torch.randn(dim, generator=g)
        sd[p + ".gamma"] = torch.full((dim,), 1.0 / n_layers) * (1 + 0.1 * torch.randn(dim, generator=g))
    sd["backbone.final_layer_norm.weight"] = 1 + 0.1 * torch.randn(dim, generator=g)
    sd["backbone.final_layer_norm.bias"] = 0.1 * torch.randn(dim, generator=g)
    sd["head.out.weight"], sd["head.out.bias"] = tn(n_fft + 2, dim) * 3, 0.1 * torch.randn(n_fft + 2, generator=g)
    sd["head.istft.window"] = torch.hann_window(n_fft)
    return sd


@torch.no_grad()
def vocos_decode(sd: dict, mel: torch.Tensor, n_fft=1024, hop=256) -> torch.Tensor:
    """mel [B, 100, T] fp32 -> wav [B, (T-1)*hop]   (Vocos.decode = ISTFTHead(VocosBackbone(mel)))."""
    sd = {k: v.to(mel.device, torch.float32) for k, v in sd.items()}
    dim = sd["backbone.embed.weight"].shape[0]
    x = F.conv1d(mel.float(), sd["backbone.embed.weight"], sd["backbone.embed.bias"], padding=3)
    x = F.layer_norm(x.transpose(1, 2), (dim,), sd["backbone.norm.weight"], sd["backbone.norm.bias"], eps=1e-6).transpose(1, 2)
    i = 0
    while f"backbone.convnext.{i}.gamma" in sd:
        p = f"backbone.convnext.{i}"
        r = x
        h = F.conv1d(x, sd[p + ".dwconv.weight"], sd[p + ".dwconv.bias"], padding=3, groups=dim).transpose(1, 2)
        h = F.layer_norm(h, (dim,), sd[p + ".norm.weight"], sd[p + ".norm.bias"], eps=1e-6)
        h = F.linear(F.gelu(F.linear(h, sd[p + ".pwconv1.weight"], sd[p + ".pwconv1.bias"])), sd[p + ".pwconv2.weight"],
                     sd[p + ".pwconv2.bias"])
        x = r + (sd[p + ".gamma"] * h).transpose(1, 2)
        i += 1
    x = F.layer_norm(x.transpose(1, 2), (dim,), sd["backbone.final_layer_norm.weight"], sd["backbone.final_layer_norm.bias"], eps=1e-6)
    x = F.linear(x, sd["head.out.weight"], sd["head.out.bias"]).transpose(1, 2)
    return istft_head(x, n_fft, hop)


def istft_head(x: torch.Tensor, n_fft=1024, hop=256) -> torch.Tensor:
    """x [B, n_fft + 2, T] = (log-magnitude | phase) -> wav; vocos ISTFTHead after its Linear."""
    mag, p = x.chunk(2, dim=1)
    mag = torch.clip(torch.exp(mag), max=1e2)
    S = mag * (torch.cos(p) + 1j * torch.sin(p))
    window = torch.hann_window(n_fft, device=x.device, dtype=x.dtype)
    return torch.istft(S, n_fft, hop, n_fft, window, center=True)


# -------------------------------------------------------------------------------------------------------
# Prompt preparation and long-form glue (TEST INFRASTRUCTURE, like the rest of this file)
# -------------------------------------------------------------------------------------------------------
def rms_floor(audio: torch.Tensor, target_rms=0.1):
    """eval/utils_eval.py:120-123 / infer/utils_infer.py:441-443."""
    rms = torch.sqrt(torch.mean(torch.square(audio)))
    if rms < target_rms:
        audio = audio * target_rms / rms
    return audio, rms


def resample(wave: torch.Tensor, orig_freq: int, new_freq: int, lowpass_filter_width=6, rolloff=0.99) -> torch.Tensor:
    """torchaudio.transforms.Resample(orig, new)(wave) (torchaudio is a third-party dependency of the reference,
    pyproject.toml; call sites eval/utils_eval.py:125-127, infer/utils_infer.py:444-446), restated from
    torchaudio.functional.functional._get_sinc_resample_kernel / _apply_sinc_resample_kernel (sinc_interp_hann).
    Pinned against torchaudio itself by tests/golden/resample.pt."""
    import math
    if orig_freq == new_freq:
        return wave
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    base_freq = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base_freq)
    idx = torch.arange(-width, width + orig, dtype=torch.float64)[None, None] / orig
    t = torch.arange(0, -new, -1, dtype=torch.float64)[:, None, None] / new + idx
    t *= base_freq
    t = t.clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t *= math.pi
    scale = base_freq / orig
    kernels = torch.where(t == 0, torch.tensor(1.0).to(t), t.sin() / t)
    kernels *= window * scale
    kernels = kernels.to(dtype=torch.float32, device=wave.device)
    shape = wave.size()
    w = wave.reshape(-1, shape[-1])
    length = w.shape[1]
    w = F.pad(w, (width, width + orig))
    res = F.conv1d(w[:, None], kernels, stride=orig).transpose(1, 2).reshape(w.shape[0], -1)
    target_length = int(math.ceil(new * length / orig))
    return res[..., :target_length].reshape(shape[:-1] + (target_length,))


def crossfade_chain(waves, cross_fade_duration=0.15, sample_rate=24000):
    """infer/utils_infer.py:521-556 (numpy; float32 chunks, float64 fades)."""
    import numpy as np
    if cross_fade_duration <= 0:
        return np.concatenate(waves)
    final = waves[0]
    for i in range(1, len(waves)):
        prev, nxt = final, waves[i]
        n = min(int(cross_fade_duration * sample_rate), len(prev), len(nxt))
        if n <= 0:
            final = np.concatenate([prev, nxt])
            continue
        mixed = prev[-n:] * np.linspace(1, 0, n) + nxt[:n] * np.linspace(0, 1, n)
        final = np.concatenate([prev[:-n], mixed, nxt[n:]])
    return final


def long_form(dit: OracleDiT, vocos_sd, ref_audio, sr, ref_text_ids, gen_text_ids, gen_text_bytes, ref_text_bytes, *,
              steps, cfg_strength=2.0, sway=-1.0, speed=1.0, target_rms=0.1, cross_fade_duration=0.15, seed=None,
              hop=256, target_sr=24000):
    """infer_batch_process (infer/utils_infer.py:418-564) with the tokenizer factored out: chunk c has token ids
    ref_text_ids + gen_text_ids[c] and UTF-8 byte counts ref_text_bytes / gen_text_bytes[c] (the duration rule uses
    bytes).  One batch-1 solve per chunk, like the reference.  Returns (final wave float64 numpy, [mel per chunk])."""
    if ref_audio.shape[0] > 1:
        ref_audio = torch.mean(ref_audio, dim=0, keepdim=True)
    audio, rms = rms_floor(ref_audio, target_rms)
    audio = resample(audio, sr, target_sr)
    ref_len = audio.shape[-1] // hop
    waves, mels = [], []
    for ids, nbytes in zip(gen_text_ids, gen_text_bytes):
        local_speed = 0.3 if nbytes < 10 else speed
        duration = ref_len + int(ref_len / ref_text_bytes * nbytes / local_speed)
        text = torch.tensor([list(ref_text_ids) + list(ids)], dtype=torch.long, device=dit.device)
        out, _ = cfm_sample(dit, audio.to(dit.device), text, duration, steps=steps, cfg_strength=cfg_strength,
                            sway_sampling_coef=sway, seed=seed)
        gen = out.to(torch.float32)[:, ref_len:, :].permute(0, 2, 1)
        wave = vocos_decode(vocos_sd, gen)
        if rms < target_rms:
            wave = wave * rms / target_rms
        waves.append(wave.squeeze().cpu().numpy())
        mels.append(gen[0].cpu().numpy())
    return crossfade_chain(waves, cross_fade_duration, target_sr), mels


def get_inference_prompt(metainfo, speed=1.0, target_sample_rate=24000, hop_length=256, target_rms=0.1,
                         infer_batch_size=1, num_buckets=200, min_secs=3, max_secs=40):
    """eval/utils_eval.py:77-219 with waveforms in place of file names and the tokenizer left out (text_list = raw
    strings): one prompt at a time on the CPU, like the reference."""
    import math
    import random
    prompts_all = []
    min_tokens = min_secs * target_sample_rate // hop_length
    max_tokens = max_secs * target_sample_rate // hop_length
    acc = [0] * num_buckets
    cols = [[[] for _ in range(num_buckets)] for _ in range(6)]

    def emit(b):
        mels = cols[2][b]
        T = max(m.shape[-1] for m in mels)
        padded = torch.stack([F.pad(m, (0, T - m.shape[-1]), value=0) for m in mels]).permute(0, 2, 1)
        prompts_all.append((cols[0][b], cols[1][b], padded, cols[3][b], cols[4][b], cols[5][b]))

    for utt, prompt_text, (ref_audio, ref_sr), gt_text, _ in metainfo:
        ref_audio = ref_audio.reshape(1, -1)
        ref_audio, ref_rms = rms_floor(ref_audio, target_rms)
        ref_audio = resample(ref_audio, ref_sr, target_sample_rate)
        if len(prompt_text[-1].encode("utf-8")) == 1:
            prompt_text = prompt_text + " "
        text_list = [prompt_text + gt_text]
        ref_mel_len = ref_audio.shape[-1] // hop_length
        total = ref_mel_len + int(ref_mel_len / len(prompt_text.encode("utf-8")) * len(gt_text.encode("utf-8")) / speed)
        ref_mel = mel_spectrogram(ref_audio).squeeze(0)
        assert min_tokens <= total <= max_tokens
        b = math.floor((total - min_tokens) / (max_tokens - min_tokens + 1) * num_buckets)
        for c, v in zip(cols, (utt, ref_rms, ref_mel, ref_mel_len, total)):
            c[b].append(v)
        cols[5][b].extend(text_list)
        acc[b] += total
        if acc[b] >= infer_batch_size:
            emit(b)
            acc[b] = 0
            for c in cols:
                c[b] = []
    for b, frames in enumerate(acc):
        if frames > 0:
            emit(b)
    random.seed(666)
    random.shuffle(prompts_all)
    return prompts_all


def eval_infer(dit: OracleDiT, vocos_sd, prompts_all, *, steps, cfg_strength=2.0, sway=-1.0, seed=None, target_rms=0.1):
    """The loop of eval/eval_infer_batch.py:187-216 (one `sample` per batch, one vocoder call per utterance, loudness
    gain); text_list entries are token-id lists here (the tokenizer is not on the path).  Returns {utt: wave [1, n]}."""
    res = {}
    for utts, ref_rms_list, ref_mels, ref_mel_lens, total_mel_lens, texts in prompts_all:
        nt = max(len(t) for t in texts)
        text = torch.full((len(texts), nt), -1, dtype=torch.long, device=dit.device)
        for i, t in enumerate(texts):
            text[i, : len(t)] = torch.tensor(t, dtype=torch.long)
        ref_len = torch.tensor(ref_mel_lens, dtype=torch.long, device=dit.device)
        total = torch.tensor(total_mel_lens, dtype=torch.long, device=dit.device)
        generated, _ = cfm_sample(dit, ref_mels.to(dit.device), text, total, lens=ref_len, steps=steps,
                                  cfg_strength=cfg_strength, sway_sampling_coef=sway, seed=seed)
        for i, gen in enumerate(generated):
            gen = gen[ref_mel_lens[i]: total_mel_lens[i], :].unsqueeze(0)
            wave = vocos_decode(vocos_sd, gen.permute(0, 2, 1).to(torch.float32)).cpu()
            if ref_rms_list[i] < target_rms:
                wave = wave * ref_rms_list[i] / target_rms
            res[utts[i]] = wave
    return res
