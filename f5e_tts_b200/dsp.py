"""MelSpec front end over libf5e_b200 (f5e_melspec).  Tables are built once per device with torch (fp64 -> fp32)."""
from __future__ import annotations

import math

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

_TABLES: dict = {}


def hz_to_mel_htk(f: float) -> float:
    return 2595.0 * math.log10(1.0 + f / 700.0)


def mel_filterbank(n_freqs=513, f_min=0.0, f_max=12000.0, n_mels=100, sample_rate=24000) -> torch.Tensor:
    """HTK triangular filterbank, norm=None, [n_freqs, n_mels] - the algorithm of torchaudio.functional.melscale_fbanks,
    which the reference reaches through torchaudio.transforms.MelSpectrogram (model/modules.py:83-93)."""
    all_freqs = torch.linspace(0, sample_rate // 2, n_freqs)
    m_pts = torch.linspace(hz_to_mel_htk(f_min), hz_to_mel_htk(f_max), n_mels + 2)
    f_pts = 700.0 * (10.0 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    down = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    return torch.clamp(torch.min(down, up), min=0.0)


def fft_tables(device):
    """(window [1024] periodic Hann fp32, twiddles float2[512] = exp(-2 pi i j / 1024))."""
    key = ("fft", str(device))
    if key not in _TABLES:
        window = torch.hann_window(1024, periodic=True, dtype=torch.float32)
        ang = -2.0 * math.pi * torch.arange(512, dtype=torch.float64) / 1024.0
        tw = torch.stack((ang.cos(), ang.sin()), dim=-1).to(torch.float32)
        _TABLES[key] = (window.to(device).contiguous(), tw.to(device).contiguous())
    return _TABLES[key]


def mel_tables(device, n_mels, sample_rate):
    key = ("mel", str(device), n_mels, sample_rate)
    if key not in _TABLES:
        fb = mel_filterbank(513, 0.0, sample_rate / 2.0, n_mels, sample_rate)  # [513, n_mels]
        nz = fb > 0
        lo = torch.where(nz.any(0), nz.float().argmax(0), torch.zeros(n_mels, dtype=torch.long))
        hi = torch.where(nz.any(0), 513 - nz.flip(0).float().argmax(0), torch.zeros(n_mels, dtype=torch.long))
        _TABLES[key] = (fb.t().contiguous().to(device), lo.to(torch.int32).to(device), hi.to(torch.int32).to(device))
    return _TABLES[key]


def mel_spectrogram(spec_module, wav: torch.Tensor) -> torch.Tensor:
    """wav [B, nw] (or [B, 1, nw]) on a CUDA device -> log-mel fp32 [B, n_mels, nw // 256 + 1]."""
    if wav.dim() == 3:
        wav = wav.squeeze(1)
    assert wav.dim() == 2
    if not wav.is_cuda:
        raise RuntimeError("f5e_tts_b200.MelSpec has no CPU path: pass a CUDA tensor")
    if (spec_module.n_fft, spec_module.hop_length, spec_module.win_length) != (1024, 256, 1024):
        raise NotImplementedError("f5e_tts_b200.MelSpec: the kernel is specialised for n_fft 1024 / hop 256 / win 1024")
    _lib.require_device()
    wav = wav.float()
    if wav.stride(1) != 1:
        wav = wav.contiguous()
    B, nw = wav.shape
    window, tw = fft_tables(wav.device)
    fbt, lo, hi = mel_tables(wav.device, spec_module.n_mel_channels, spec_module.target_sample_rate)
    out = torch.empty(B, spec_module.n_mel_channels, nw // 256 + 1, device=wav.device, dtype=torch.float32)
    check(_lib.load().f5e_melspec(ptr(wav), _lib.C.c_int64(wav.stride(0)), B, nw, ptr(window), ptr(tw), ptr(fbt), ptr(lo), ptr(hi),
                                  spec_module.n_mel_channels, ptr(out), stream_ptr()), "f5e_melspec")
    return out
