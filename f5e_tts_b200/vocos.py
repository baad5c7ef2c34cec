"""Vocos (mel-24khz) vocoder mirror backed by libf5e_b200 (f5e_vocos_decode).

The reference takes its vocoder from the third-party `vocos` package (`Vocos.from_hparams(config.yaml)` +
`load_state_dict(pytorch_model.bin)`, infer/utils_infer.py:101-124) and only ever calls `vocoder.decode(mel)`
(utils_infer.py:488-489, eval/eval_infer_batch.py:211-212).  This module keeps the package's parameter names
(`backbone.embed`, `backbone.convnext.N.{dwconv,norm,pwconv1,pwconv2,gamma}`, `backbone.final_layer_norm`, `head.out`,
`head.istft.window`, `feature_extractor.mel_spec.*`) so the published checkpoint loads unchanged.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from . import _lib
from ._lib import check, ptr, stream_ptr


class _Block(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("dw_w", "dw_b", "ln_w", "ln_b", "w1", "b1", "w2", "b2", "gamma")]


class _Model(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("dim", "inter", "n_layers", "in_ch", "k_embed", "n_fft", "hop")] + [
        (n, C.c_void_p) for n in ("w_embed", "b_embed", "ln0_w", "ln0_b")] + [("blocks", C.POINTER(_Block))] + [
        (n, C.c_void_p) for n in ("lnf_w", "lnf_b", "w_head", "b_head", "window", "twiddles")]


class _Buffers(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("a_embed", "x", "h", "a1", "a2", "spec")]


def split3_weight(w: torch.Tensor) -> torch.Tensor:
    """fp32 [N, K] -> bf16 [N, 3K] = [W_hi | W_lo | W_hi] (pairs with activations split as [A_hi | A_hi | A_lo])."""
    w = w.float()
    hi = w.to(torch.bfloat16)
    lo = (w - hi.float()).to(torch.bfloat16)
    return torch.cat([hi, lo, hi], dim=1).contiguous()


class _ConvNeXtBlock(nn.Module):
    def __init__(self, dim, inter, layer_scale):
        super().__init__()
        self.dwconv = nn.Conv1d(dim, dim, 7, padding=3, groups=dim)
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.pwconv1 = nn.Linear(dim, inter)
        self.pwconv2 = nn.Linear(inter, dim)
        self.gamma = nn.Parameter(layer_scale * torch.ones(dim))


class _Backbone(nn.Module):
    def __init__(self, in_ch, dim, inter, n_layers):
        super().__init__()
        self.embed = nn.Conv1d(in_ch, dim, 7, padding=3)
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.convnext = nn.ModuleList([_ConvNeXtBlock(dim, inter, 1.0 / n_layers) for _ in range(n_layers)])
        self.final_layer_norm = nn.LayerNorm(dim, eps=1e-6)


class _ISTFT(nn.Module):
    def __init__(self, n_fft):
        super().__init__()
        self.register_buffer("window", torch.hann_window(n_fft))


class _Head(nn.Module):
    def __init__(self, dim, n_fft):
        super().__init__()
        self.out = nn.Linear(dim, n_fft + 2)
        self.istft = _ISTFT(n_fft)


class Vocos(nn.Module):
    """charactr/vocos-mel-24khz architecture: VocosBackbone(100, 512, 1536, 8) + ISTFTHead(512, 1024, 256, "center")."""

    def __init__(self, input_channels=100, dim=512, intermediate_dim=1536, num_layers=8, n_fft=1024, hop_length=256, seed=None):
        super().__init__()
        if seed is not None:
            torch.manual_seed(seed)
        self.input_channels, self.dim, self.intermediate_dim, self.num_layers = input_channels, dim, intermediate_dim, num_layers
        self.n_fft, self.hop_length = n_fft, hop_length
        self.backbone = _Backbone(input_channels, dim, intermediate_dim, num_layers)
        self.head = _Head(dim, n_fft)
        for m in self.modules():  # the package's init: trunc_normal_(std=0.02) weights, zero biases
            if isinstance(m, (nn.Conv1d, nn.Linear)):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.constant_(m.bias, 0)
        self._plan = None
        self._ws = {}

    @classmethod
    def from_hparams(cls, config_path: str) -> "Vocos":
        import yaml

        cfg = yaml.safe_load(open(config_path))
        bb, hd = cfg["backbone"]["init_args"], cfg["head"]["init_args"]
        return cls(bb["input_channels"], bb["dim"], bb["intermediate_dim"], bb["num_layers"], hd["n_fft"], hd["hop_length"])

    def load_state_dict(self, state_dict, strict=True, **k):
        self._plan, self._ws = None, {}
        sd = {k_: v for k_, v in state_dict.items() if not k_.startswith("feature_extractor.")}
        return super().load_state_dict(sd, strict=strict, **k)

    def _apply(self, fn, *a, **k):
        self._plan, self._ws = None, {}
        return super()._apply(fn, *a, **k)

    def _load_from_state_dict(self, *a, **k):  # reached when a PARENT module's load_state_dict loads this one
        self._plan, self._ws = None, {}
        return super()._load_from_state_dict(*a, **k)

    def _param_versions(self):  # in-place parameter updates (copy_, optimizer steps) bump these
        return tuple(p._version for p in self.parameters())

    def forward(self, *a, **k):
        raise NotImplementedError("f5e_tts_b200.Vocos is decode-only (the reference never calls vocoder.forward)")

    def _get_plan(self):
        if self._plan is not None:
            if self._plan[-1] == self._param_versions():
                return self._plan
            self._plan, self._ws = None, {}
        dev = self.head.out.weight.device
        if dev.type != "cuda":
            raise RuntimeError("f5e_tts_b200.Vocos has no CPU path: move the module to a B200 (cuda) device")
        _lib.require_device()
        from .dsp import fft_tables

        keep = []

        def F(t):
            t = t.detach().to(dev, torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        def W3(t):
            t = split3_weight(t.detach().to(dev))
            keep.append(t)
            return t.data_ptr()

        bb = self.backbone
        m = _Model()
        m.dim, m.inter, m.n_layers, m.in_ch = self.dim, self.intermediate_dim, self.num_layers, self.input_channels
        m.k_embed = (7 * self.input_channels + 63) // 64 * 64
        m.n_fft, m.hop = self.n_fft, self.hop_length
        we = torch.zeros(self.dim, m.k_embed, device=dev)
        we[:, : 7 * self.input_channels] = bb.embed.weight.detach().float().permute(0, 2, 1).reshape(self.dim, -1)  # [co, tap*C + c]
        m.w_embed, m.b_embed = W3(we), F(bb.embed.bias)
        m.ln0_w, m.ln0_b = F(bb.norm.weight), F(bb.norm.bias)
        blocks = (_Block * self.num_layers)()
        for i, b in enumerate(bb.convnext):
            blocks[i].dw_w, blocks[i].dw_b = F(b.dwconv.weight.reshape(self.dim, 7)), F(b.dwconv.bias)
            blocks[i].ln_w, blocks[i].ln_b = F(b.norm.weight), F(b.norm.bias)
            blocks[i].w1, blocks[i].b1 = W3(b.pwconv1.weight), F(b.pwconv1.bias)
            blocks[i].w2, blocks[i].b2 = W3(b.pwconv2.weight), F(b.pwconv2.bias)
            blocks[i].gamma = F(b.gamma)
        m.blocks = C.cast(blocks, C.POINTER(_Block))
        m.lnf_w, m.lnf_b = F(bb.final_layer_norm.weight), F(bb.final_layer_norm.bias)
        m.w_head, m.b_head = W3(self.head.out.weight), F(self.head.out.bias)
        window, tw = fft_tables(dev)
        m.window, m.twiddles = window.data_ptr(), tw.data_ptr()
        self._plan = (m, blocks, keep, (window, tw), self._param_versions())
        return self._plan

    def _workspace(self, batch, t_pad):
        key = (batch, t_pad)
        if key not in self._ws:
            dev, rows = self.head.out.weight.device, batch * t_pad
            m = self._get_plan()[0]
            t = dict(a_embed=torch.empty(rows, 3 * m.k_embed, device=dev, dtype=torch.bfloat16),
                     x=torch.empty(rows, self.dim, device=dev), h=torch.empty(rows, self.dim, device=dev),
                     a1=torch.empty(rows, 3 * self.dim, device=dev, dtype=torch.bfloat16),
                     a2=torch.empty(rows, 3 * self.intermediate_dim, device=dev, dtype=torch.bfloat16),
                     spec=torch.empty(rows, (self.n_fft + 2 + 3) // 4 * 4, device=dev))
            self._ws = {key: (t, _Buffers(**{k: v.data_ptr() for k, v in t.items()}))}
        return self._ws[key]

    def _decode(self, mel, ld_b, ld_c, ld_t, t_off, n_frames, batch, t_pad, out_len):
        m = self._get_plan()[0]
        _, buf = self._workspace(batch, t_pad)
        out = torch.empty(batch, max(out_len, 0), device=mel.device, dtype=torch.float32)
        if out_len <= 0:
            return out
        check(_lib.load().f5e_vocos_decode(C.byref(m), C.byref(buf), ptr(mel), C.c_int64(ld_b), C.c_int64(ld_c), C.c_int64(ld_t),
                                           ptr(t_off), ptr(n_frames), batch, t_pad, ptr(out), C.c_int64(out.stride(0)), out_len,
                                           stream_ptr()), "f5e_vocos_decode")
        return out

    @torch.no_grad()
    def decode(self, features_input: torch.Tensor) -> torch.Tensor:
        """mel [B, C, T] float32 -> waveform [B, (T - 1) * hop] (same contract as vocos.Vocos.decode)."""
        x = features_input
        if not x.is_cuda:
            raise RuntimeError("f5e_tts_b200.Vocos.decode has no CPU path: pass a CUDA tensor")
        if x.dtype != torch.float32:
            x = x.float()
        B, Cc, T = x.shape
        assert Cc == self.input_channels
        return self._decode(x, x.stride(0), x.stride(1), x.stride(2), None, None, B, T, (T - 1) * self.hop_length)

    @torch.no_grad()
    def decode_varlen(self, mel_btc: torch.Tensor, start: torch.Tensor, end: torch.Tensor, max_frames: int | None = None) -> torch.Tensor:
        """Batched form of the reference's per-utterance loop (eval/eval_infer_batch.py:208-212):
        for every sample b decode mel_btc[b, start[b]:end[b], :] (layout [B, T, C], any float dtype) exactly as if it were
        decoded alone.  Returns fp32 [B, (max_frames - 1) * hop], zero padded; sample b is valid up to
        (end[b] - start[b] - 1) * hop.  `max_frames` avoids a device->host sync when the caller knows it."""
        x = mel_btc
        if x.dtype != torch.float32:
            x = x.float()
        B = x.shape[0]
        t_off = start.to(device=x.device, dtype=torch.int32)
        n_fr = (end.to(x.device) - start.to(x.device)).to(torch.int32)
        if max_frames is None:
            max_frames = int(n_fr.max())
        return self._decode(x, x.stride(0), x.stride(2), x.stride(1), t_off, n_fr, B, max_frames, (max_frames - 1) * self.hop_length)
