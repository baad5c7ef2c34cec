"""PPG front end on the GPU: f5e_kaldi_fbank (+ fused GlobalCMVN) against the oracle and against torchaudio's kaldi.fbank."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import ppg_oracle as PO  # noqa: E402

pytestmark = pytest.mark.gpu

# fp32 log-mel energies.  Pre-emphasis leaves the low mel bins ~600x weaker than the rest, so fp32 FFT round-off (relative to
# the LARGEST bins) shows there: torchaudio's own fp32 result is 1.3e-4 away from a float64 evaluation on these inputs, the
# kernel (radix-2, two frames per FFT) 3.3e-4 away from torchaudio.  Bar: 1e-3 absolute on the log energies (0.1 % energy).
TOL = 1e-3


@pytest.mark.parametrize("n", [400, 1999, 16000, 160000])
def test_fbank_vs_oracle(cuda_dev, n):
    from f5e_tts_b200.ppg import kaldiFbank
    g = torch.Generator().manual_seed(n)
    wave = 0.1 * torch.randn(2, n, generator=g)
    wave[1] *= torch.linspace(0.0, 1.0, n)  # a quiet start: small energies
    got, lens = kaldiFbank()(wave.to(cuda_dev))
    assert lens.tolist() == [1 + (n - 400) // 160] and got.shape == (2, int(lens[0]), 80)
    for b in range(2):
        want = PO.kaldi_fbank(wave[b])
        err = (got[b].cpu() - want).abs().max().item()
        print(f"n={n} utt {b}: max |log-mel diff| {err:.2e}")
        assert err < TOL


def test_fbank_vs_torchaudio_and_cmvn(cuda_dev):
    kaldi = pytest.importorskip("torchaudio.compliance.kaldi")
    from f5e_tts_b200.ppg import GlobalCMVN, kaldiFbank
    g = torch.Generator().manual_seed(7)
    wave = 0.2 * torch.randn(1, 32000, generator=g)
    want = kaldi.fbank(wave * (1 << 15), num_mel_bins=80, frame_length=25, frame_shift=10, dither=0.0, energy_floor=0.0,
                       sample_frequency=16000)
    got, _ = kaldiFbank()(wave.to(cuda_dev))
    assert (got[0].cpu() - want).abs().max().item() < TOL
    mean, istd = torch.randn(80, generator=g) + 10.0, torch.rand(80, generator=g) + 0.5
    cm = GlobalCMVN(mean, istd)
    got_c, _ = kaldiFbank()(wave.to(cuda_dev), cmvn=cm)
    assert torch.equal(got_c[0].cpu(), PO.global_cmvn(got[0].cpu(), mean, istd))  # same two fp32 roundings


def test_fbank_ragged_batch_equals_single_calls(cuda_dev):
    from f5e_tts_b200.ppg import kaldiFbank
    g = torch.Generator().manual_seed(11)
    lens = torch.tensor([24000, 399, 8000, 15999])
    wave = 0.1 * torch.randn(4, 24000, generator=g)
    fb = kaldiFbank()
    got, n_frames = fb(wave.to(cuda_dev), lens=lens.to(cuda_dev))
    assert n_frames.tolist() == [148, 0, 48, 98]
    for b in range(4):
        T = int(n_frames[b])
        if T:
            alone, _ = fb(wave[b:b + 1, :int(lens[b])].to(cuda_dev))
            assert torch.equal(got[b, :T], alone[0])
        assert float(got[b, T:].abs().max()) == 0.0 if T < got.shape[1] else True


def test_fbank_rejects_cpu_and_short(cuda_dev):
    from f5e_tts_b200.ppg import kaldiFbank
    with pytest.raises(RuntimeError):
        kaldiFbank()(torch.zeros(1, 1600))
    with pytest.raises(ValueError):
        kaldiFbank()(torch.zeros(1, 399, device=cuda_dev))
