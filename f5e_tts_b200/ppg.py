"""Front end of the PPG extractor on the GPU (first half): the reference's `kaldiFbank`
(ppg/wenet/dataset/feats.py:49-83) and `GlobalCMVN` (ppg/wenet/transformer/cmvn.py:9-47) as ONE kernel
(`f5e_kaldi_fbank`), plus the length / padding-mask logic of `PPGModelWapper` (ppg/ppg_model.py:110-168).

Same names and argument meaning as the reference.  `kaldiFbank.forward` additionally takes `lens` (a padded batch of
utterances of different length: every utterance gets the frames a batch-1 call would give it, the rest are zeros) and
`cmvn` (fuses the encoder's GlobalCMVN into the same pass).  There is no CPU path.
"""
from __future__ import annotations

import math

import torch
from torch import nn

from . import _lib
from ._lib import check, ptr, stream_ptr

_TABLES: dict = {}


def povey_window(win_length: int) -> torch.Tensor:
    """torchaudio.compliance.kaldi window_type='povey': hann(periodic=False) ** 0.85, fp32."""
    return torch.hann_window(win_length, periodic=False, dtype=torch.float32).pow(0.85)


def kaldi_mel_banks(num_bins=80, padded=512, sample_freq=16000.0, low_freq=20.0, high_freq=0.0) -> torch.Tensor:
    """fp32 [num_bins, padded // 2 + 1]: the algorithm of torchaudio.compliance.kaldi.get_mel_banks (no VTLN) with the
    zero Nyquist column kaldi.fbank appends; same torch fp32 operations in the same order, so the table matches bit for bit."""
    num_fft_bins = padded // 2
    nyquist = 0.5 * sample_freq
    if high_freq <= 0.0:
        high_freq += nyquist
    fft_bin_width = sample_freq / padded
    mel = lambda f: 1127.0 * math.log(1.0 + f / 700.0)
    mel_low, mel_high = mel(low_freq), mel(high_freq)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = torch.arange(num_bins).unsqueeze(1)
    left, center, right = mel_low + b * delta, mel_low + (b + 1.0) * delta, mel_low + (b + 2.0) * delta
    m = (1127.0 * (1.0 + fft_bin_width * torch.arange(num_fft_bins) / 700.0).log()).unsqueeze(0)
    up, down = (m - left) / (center - left), (right - m) / (right - center)
    bins = torch.max(torch.zeros(1), torch.min(up, down))
    return torch.nn.functional.pad(bins, (0, 1), mode="constant", value=0.0).float().contiguous()


def _fbank_tables(device, n_mels, win, n_fft, sample_rate):
    key = (str(device), n_mels, win, n_fft, sample_rate)
    if key not in _TABLES:
        fb = kaldi_mel_banks(n_mels, n_fft, float(sample_rate))
        nz = fb > 0
        n_bin = fb.shape[1]
        lo = torch.where(nz.any(1), nz.float().argmax(1), torch.zeros(n_mels, dtype=torch.long))
        hi = torch.where(nz.any(1), n_bin - nz.flip(1).float().argmax(1), torch.zeros(n_mels, dtype=torch.long))
        ang = -2.0 * math.pi * torch.arange(n_fft // 2, dtype=torch.float64) / n_fft
        tw = torch.stack((ang.cos(), ang.sin()), dim=-1).to(torch.float32)
        _TABLES[key] = tuple(t.to(device).contiguous() for t in
                             (povey_window(win), tw, fb, lo.to(torch.int32), hi.to(torch.int32)))
    return _TABLES[key]


class GlobalCMVN(nn.Module):
    """ppg/wenet/transformer/cmvn.py:9-47: (x - mean) * istd.  Used fused (`kaldiFbank.forward(cmvn=...)`)."""

    def __init__(self, mean: torch.Tensor, istd: torch.Tensor, norm_var: bool = True):
        super().__init__()
        assert mean.shape == istd.shape
        self.norm_var = norm_var
        self.register_buffer("mean", mean.float())
        self.register_buffer("istd", istd.float())


class kaldiFbank(nn.Module):
    """ppg/wenet/dataset/feats.py:49-83.  forward(x [B, L] in [-1, 1]) -> (feats fp32 [B, T, n_mels], speech_lengths)."""

    def __init__(self, sample_rate=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80, spec_mask_time=(5, 10),
                 spec_mask_freq=(5, 10)):
        super().__init__()
        self.f_length = int(win_length / sample_rate * 1000)  # ms, as the reference passes them on
        self.f_shift = int(hop_length / sample_rate * 1000)
        self.sample_rate, self.n_fft, self.n_mels = sample_rate, n_fft, n_mels
        if n_fft != 512:
            raise NotImplementedError("f5e_tts_b200.kaldiFbank: the kernel is specialised for the 512-point FFT")

    @torch.no_grad()
    def forward(self, x: torch.Tensor, is_spec_aug=(), *, lens: torch.Tensor | None = None, cmvn: GlobalCMVN | None = None):
        if len(is_spec_aug) != 0:
            raise NotImplementedError("f5e_tts_b200.kaldiFbank: SpecAugment is a training-time option (inference path only)")
        if not x.is_cuda:
            raise RuntimeError("f5e_tts_b200.kaldiFbank has no CPU path: pass a CUDA tensor")
        assert x.dim() == 2
        _lib.require_device()
        x = x.float()
        if x.stride(1) != 1:
            x = x.contiguous()
        B, L = x.shape
        # kaldi: window_size = int(sample_frequency * frame_length * 0.001), padded to the next power of two
        win, shift = int(self.sample_rate * self.f_length * 0.001), int(self.sample_rate * self.f_shift * 0.001)
        if L < win:
            raise ValueError(f"kaldiFbank: {L} samples are shorter than one {win}-sample frame")
        t_max = 1 + (L - win) // shift
        window, tw, fb, lo, hi = _fbank_tables(x.device, self.n_mels, win, self.n_fft, self.sample_rate)
        lens32 = None if lens is None else lens.to(device=x.device, dtype=torch.int32).contiguous()
        out = torch.empty(B, t_max, self.n_mels, device=x.device, dtype=torch.float32)
        mean = istd = None
        if cmvn is not None:
            mean = cmvn.mean.to(x.device).contiguous()
            istd = (cmvn.istd if cmvn.norm_var else torch.ones_like(cmvn.istd)).to(x.device).contiguous()
        f32, i64 = _lib.C.c_float, _lib.C.c_int64
        check(_lib.load().f5e_kaldi_fbank(ptr(x), i64(x.stride(0)), ptr(lens32), B, L, f32(float(1 << 15)), f32(0.97), win, shift,
                                          ptr(window), ptr(tw), ptr(fb), ptr(lo), ptr(hi), self.n_mels, ptr(mean), ptr(istd),
                                          ptr(out), i64(out.stride(0)), t_max, stream_ptr()), "f5e_kaldi_fbank")
        if lens is None:
            speech_lengths = torch.tensor([t_max])  # feats.py:82: one length for the whole (equal-length) batch
        else:
            n = lens.to(torch.long)
            speech_lengths = torch.where(n >= win, 1 + (n - win) // shift, torch.zeros_like(n))
        return out, speech_lengths


def make_pad_mask(lengths: torch.Tensor, max_len: int = 0) -> torch.Tensor:
    """ppg/ppg_model.py:31-57: True on the padded part."""
    max_len = max_len if max_len > 0 else int(lengths.max())
    return torch.arange(max_len, device=lengths.device).unsqueeze(0) >= lengths.unsqueeze(-1)


def ppg_true_len(mel_lens: torch.Tensor, ppg_len: int, ppg_frame_length=20, mel_f_shift=10) -> torch.Tensor:
    """ppg/ppg_model.py:139-141: frames of the (sub-sampled) PPG that belong to each utterance."""
    return (mel_lens / (ppg_frame_length / mel_f_shift)).long().clamp(max=ppg_len)
