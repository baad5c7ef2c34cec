"""CPU/torch restatement of the PPG extractor's front end.  TEST INFRASTRUCTURE ONLY.

Follows, by behaviour, /root/reference/src/f5_tts/ppg/: wenet/dataset/feats.py:49-83 (kaldiFbank),
wenet/transformer/cmvn.py:9-47 (GlobalCMVN), wenet/transformer/{subsampling.py:68-119, embedding.py:19-111,
attention.py:16-222, convolution.py:16-134, positionwise_feed_forward.py, encoder_layer.py:118-268, encoder.py:33-208,
403-495} (Conformer encoder, non-streaming), asr_model.py:222-244 (extract) and ppg_model.py:110-145 (lengths, mask).

Pinning: kaldi.fbank lives in torchaudio (third party, installed in this image): `kaldi_fbank` below restates its
published algorithm and tests/test_ppg_cpu.py checks it against torchaudio.compliance.kaldi.fbank itself.  The encoder
restatement is pinned against the reference's own wenet modules imported in the build container with random weights
(oracle/make_golden.py -> tests/golden/ppg_conformer.pt).  The pretrained PPG checkpoint, its train.yaml and global_cmvn
are NOT in the reference tree (ppg_model.py:11-29 loads them from outside), so the encoder configuration used here is
the code's defaults with a 12-block stack ("PPG_20ms model with 12 layers", ppg_model.py:12): parity on the real
checkpoint is unpinned.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

EPS = torch.finfo(torch.float32).eps


def povey_window(win: int) -> torch.Tensor:
    return torch.hann_window(win, periodic=False, dtype=torch.float32).pow(0.85)


def kaldi_mel_banks(num_bins=80, padded=512, sample_freq=16000.0, low_freq=20.0, high_freq=0.0) -> torch.Tensor:
    """[num_bins, padded/2] triangular banks on the kaldi mel scale (1127 ln(1 + f/700)), as kaldi.get_mel_banks without VTLN."""
    nyq = 0.5 * sample_freq
    if high_freq <= 0.0:
        high_freq += nyq
    mel = lambda f: 1127.0 * math.log(1.0 + f / 700.0)
    lo, hi = mel(low_freq), mel(high_freq)
    delta = (hi - lo) / (num_bins + 1)
    b = torch.arange(num_bins).unsqueeze(1)
    left, center, right = lo + b * delta, lo + (b + 1.0) * delta, lo + (b + 2.0) * delta
    m = (1127.0 * (1.0 + (sample_freq / padded) * torch.arange(padded // 2) / 700.0).log()).unsqueeze(0)
    return torch.max(torch.zeros(1), torch.min((m - left) / (center - left), (right - m) / (right - center)))


def kaldi_fbank(wave: torch.Tensor, sample_rate=16000, n_mels=80, frame_length_ms=25, frame_shift_ms=10,
                preemph=0.97) -> torch.Tensor:
    """feats.py:59-71 for ONE utterance: wave [L] in [-1, 1] -> [T, n_mels] log mel energies.
    kaldi.fbank defaults: snip_edges, remove_dc_offset, povey window, round_to_power_of_two, use_power, use_log_fbank."""
    x = wave.float() * (1 << 15)
    win, shift = int(sample_rate * frame_length_ms * 0.001), int(sample_rate * frame_shift_ms * 0.001)
    padded = 1 << (win - 1).bit_length()
    if x.numel() < win:
        return x.new_zeros(0, n_mels)
    T = 1 + (x.numel() - win) // shift
    fr = x.as_strided((T, win), (shift, 1)).clone()
    fr = fr - fr.mean(dim=1, keepdim=True)
    prev = torch.cat([fr[:, :1], fr[:, :-1]], dim=1)  # replicate padding on the left
    fr = fr - preemph * prev
    fr = fr * povey_window(win).to(fr.device)
    fr = F.pad(fr, (0, padded - win))
    power = torch.fft.rfft(fr).abs().pow(2.0)
    banks = F.pad(kaldi_mel_banks(n_mels, padded, float(sample_rate)), (0, 1)).to(fr.device)
    return torch.max(power @ banks.T, torch.tensor(EPS, device=fr.device)).log()


def global_cmvn(x: torch.Tensor, mean: torch.Tensor, istd: torch.Tensor) -> torch.Tensor:
    """cmvn.py:34-47."""
    return (x - mean) * istd
