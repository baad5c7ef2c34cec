"""MelSpec front end (mirror of reference model/modules.py:104-143).  The CUDA STFT/mel kernel (csrc/dsp.cu) is bound
in `MelSpec.forward` through dsp.mel_spectrogram."""
from __future__ import annotations

import torch
from torch import nn


class MelSpec(nn.Module):
    def __init__(self, n_fft=1024, hop_length=256, win_length=1024, n_mel_channels=100, target_sample_rate=24_000,
                 mel_spec_type="vocos"):
        super().__init__()
        if mel_spec_type != "vocos":
            raise NotImplementedError("f5e_tts_b200.MelSpec: only the vocos mel front end is implemented (bigvgan is out of scope)")
        self.n_fft, self.hop_length, self.win_length = n_fft, hop_length, win_length
        self.n_mel_channels, self.target_sample_rate = n_mel_channels, target_sample_rate
        self.register_buffer("dummy", torch.tensor(0), persistent=False)

    def forward(self, wav):
        from . import dsp
        if self.dummy.device != wav.device:  # the reference moves itself to the input's device (modules.py:131-132)
            self.to(wav.device)
        return dsp.mel_spectrogram(self, wav)
