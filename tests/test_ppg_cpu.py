"""PPG front end, CPU side: the oracle's kaldi fbank restatement against torchaudio.compliance.kaldi.fbank itself (the
third-party implementation the reference calls, ppg/wenet/dataset/feats.py:59-71) and the product's host tables."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import ppg_oracle as PO  # noqa: E402

kaldi = pytest.importorskip("torchaudio.compliance.kaldi")


def ref_fbank(wave):
    return kaldi.fbank((wave * (1 << 15)).unsqueeze(0), num_mel_bins=80, frame_length=25, frame_shift=10, dither=0.0,
                       energy_floor=0.0, sample_frequency=16000)


@pytest.mark.parametrize("n", [400, 559, 560, 16000, 48123])
def test_oracle_fbank_matches_torchaudio(n):
    g = torch.Generator().manual_seed(n)
    wave = 0.1 * torch.randn(n, generator=g)
    want, got = ref_fbank(wave), PO.kaldi_fbank(wave)
    assert got.shape == want.shape == (1 + (n - 400) // 160, 80)
    assert (got - want).abs().max().item() < 1e-4


def test_oracle_fbank_short_and_silent():
    assert PO.kaldi_fbank(torch.zeros(399)).shape == (0, 80)
    got, want = PO.kaldi_fbank(torch.zeros(1600)), ref_fbank(torch.zeros(1600))
    assert torch.equal(got, want)  # log(eps) everywhere


def test_product_tables_are_torchaudio_tables():
    from f5e_tts_b200.ppg import kaldi_mel_banks, povey_window
    fb, _ = kaldi.get_mel_banks(80, 512, 16000.0, 20.0, 0.0, 100.0, -500.0, 1.0)
    mine = kaldi_mel_banks()
    assert torch.equal(mine[:, :256], fb) and float(mine[:, 256].abs().max()) == 0.0
    assert torch.equal(povey_window(400), kaldi._feature_window_function("povey", 400, 0.42, torch.device("cpu"), torch.float32))


def test_lengths_and_mask():
    from f5e_tts_b200.ppg import make_pad_mask, ppg_true_len
    lens = torch.tensor([5, 3, 2])
    assert make_pad_mask(lens).tolist() == [[False] * 5, [False] * 3 + [True] * 2, [False] * 2 + [True] * 3]
    assert ppg_true_len(torch.tensor([98, 99, 301]), 120).tolist() == [49, 49, 120]


def test_oracle_conformer_matches_reference_golden():
    """tests/golden/ppg_conformer.pt: the reference's kaldiFbank + wenet ConformerEncoder + content Linear, run in the
    build container by oracle/make_golden.py (case_ppg_conformer) with the same seeded weights."""
    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", "ppg_conformer.pt"))
    cfg = PO.ConformerConfig(blocks=fx["blocks"])
    sd = PO.make_conformer_state_dict(cfg, seed=fx["seed"])
    feats = PO.kaldi_fbank(fx["wave"])
    assert (feats - fx["feats"]).abs().max().item() < 1e-4
    got = PO.conformer_extract(cfg, sd, fx["feats"])
    assert got.shape == fx["ppg"].shape == ((feats.shape[0] - 1) // 2, cfg.dim)
    rel = ((got - fx["ppg"]).norm() / fx["ppg"].norm()).item()
    assert rel < 1e-5, rel
    ppg, n = PO.audio_to_ppg(cfg, sd, fx["wave"])
    assert n == got.shape[0] and ppg.shape == got.shape


def test_oracle_conformer_causal_layernorm_matches_reference_golden():
    """The streaming recipe's conv module (causal, LayerNorm): the causal left context is glu(bias), not zero
    (convolution.py:104-113 pads before the biased pointwise conv) - the reference golden pins that."""
    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", "ppg_conformer_causal_ln.pt"))
    cfg = PO.ConformerConfig(blocks=fx["blocks"], causal=fx["causal"], cnn_norm=fx["cnn_norm"])
    sd = PO.make_conformer_state_dict(cfg, seed=fx["seed"])
    got = PO.conformer_extract(cfg, sd, fx["feats"])
    rel = ((got - fx["ppg"]).norm() / fx["ppg"].norm()).item()
    assert rel < 1e-5, rel


def test_ppg_map_golden():
    """output_type "map" (ppg_model.py:108-127) of the oracle against the reference's own PPGModelWapper.ppg_to_target."""
    import os

    import torch

    from oracle import ppg_oracle as PO

    d = torch.load(os.path.join(os.path.dirname(__file__), "golden", "ppg_map.pt"), weights_only=False)
    for ratio, key in ((1.0, "out_1"), (0.35, "out_035")):
        got = PO.ppg_to_target(d["ppg"], d["true_len"], "map", d["w"], d["b"], d["center"], ratio)
        assert torch.allclose(got, d[key], atol=1e-6, rtol=1e-6)
        assert (got[1, 17:] == 0).all()
