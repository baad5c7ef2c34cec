"""GPU parity of the prompt-preparation / long-form pipeline (f5e_tts_b200.pipeline, through the C ABI) against the oracle
restatements and the fixtures made from the reference's own code."""
import os

import numpy as np
import pytest
import torch

from oracle import f5e_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def test_rms_normalize(cuda_dev):
    from f5e_tts_b200 import pipeline as P
    g = torch.Generator().manual_seed(1)
    lens = [70001, 5003, 33000]
    amp = [0.01, 0.5, 0.099]
    wav = torch.zeros(3, max(lens))
    for i, (n, a) in enumerate(zip(lens, amp)):
        wav[i, :n] = a * torch.randn(n, generator=g)
    out, rms = P.rms_normalize(wav.to(cuda_dev), lens, 0.1)
    out2, rms2 = P.rms_normalize(wav.to(cuda_dev), lens, 0.1)
    assert torch.equal(out, out2) and torch.equal(rms, rms2)  # fixed reduction order
    for i, n in enumerate(lens):
        ref, ref_rms = O.rms_floor(wav[i:i + 1, :n], 0.1)
        assert abs(rms[i].item() - ref_rms.item()) <= 2e-6 * ref_rms.item()
        assert torch.allclose(out[i, :n].cpu(), ref[0], rtol=3e-6, atol=0)
        assert out[i, n:].abs().max().item() == 0 if n < max(lens) else True
        assert (amp[i] >= 0.12) == torch.equal(out[i, :n].cpu(), wav[i, :n])  # loud prompts pass through untouched


def test_resample_matches_torchaudio_fixture(cuda_dev):
    from f5e_tts_b200 import pipeline as P
    for c in torch.load(os.path.join(G, "resample.pt"), weights_only=False):
        y, lens = P.resample(c["x"].to(cuda_dev), c["orig"], c["new"])
        assert y.shape == c["y"].shape and int(lens[0]) == c["y"].shape[1]
        assert (y.cpu() - c["y"]).abs().max().item() < 5e-6, (c["orig"], c["new"])
        # ragged rows: a shorter row equals resampling that row alone
        n1 = c["x"].shape[1] - 1234
        y2, lens2 = P.resample(c["x"].to(cuda_dev), c["orig"], c["new"], lens=[c["x"].shape[1], n1])
        alone = O.resample(c["x"][1:2, :n1], c["orig"], c["new"])
        assert int(lens2[1]) == alone.shape[1]
        assert (y2[1, : alone.shape[1]].cpu() - alone[0]).abs().max().item() < 5e-6
        assert y2[1, alone.shape[1]:].abs().max().item() == 0


@pytest.mark.parametrize("lens,cf,gain", [([9000, 5000, 2000, 12000], 0.15, None), ([4000, 100, 7000], 0.15, (0.031, 0.1)),
                                          ([5000], 0.15, None), ([3000, 3000], 0.0, (0.05, 0.1)), ([3000, 1, 3000], 0.15, None)])
def test_crossfade_bit_exact_vs_numpy_chain(cuda_dev, lens, cf, gain):
    from f5e_tts_b200 import pipeline as P
    rng = np.random.default_rng(3)
    waves = [rng.standard_normal(n).astype(np.float32) for n in lens]
    pad = torch.zeros(len(lens), max(lens))
    for i, w in enumerate(waves):
        pad[i, : len(w)] = torch.from_numpy(w)
    got = P.crossfade_chunks(pad.to(cuda_dev), lens, cf, 24000, gain).cpu().numpy()
    if gain is not None:  # `generated_wave * rms / target_rms` in fp32 torch (utils_infer.py:495-496)
        waves = [(torch.from_numpy(w) * torch.tensor(gain[0]) / gain[1]).numpy() for w in waves]
    ref = O.crossfade_chain(waves, cf, 24000) if cf > 0 else np.concatenate(waves).astype(np.float64)
    assert got.dtype == np.float64 and got.shape == ref.shape
    assert np.array_equal(got, ref)


def _metainfo(n, seed):
    g = torch.Generator().manual_seed(seed)
    rows = []
    for i in range(n):
        sr = [16000, 24000, 44100][i % 3]
        secs = 2.0 + 3.0 * torch.rand(1, generator=g).item()
        wave = (0.02 + 0.2 * torch.rand(1, generator=g).item()) * torch.randn(1, int(secs * sr), generator=g)
        ptxt = "prompt text number %d, of some length." % i
        gtxt = " " + "generated words here, " * int(1 + 3 * torch.rand(1, generator=g).item())
        rows.append((f"utt{i}", ptxt, (wave, sr), gtxt, None))
    return rows


def test_get_inference_prompt_vs_oracle(cuda_dev):
    from f5e_tts_b200 import pipeline as P
    meta = _metainfo(14, 5)
    kw = dict(infer_batch_size=2500, num_buckets=20, min_secs=1, max_secs=40)
    got = P.get_inference_prompt(meta, device=cuda_dev, slab=4, **kw)
    ref = O.get_inference_prompt(meta, **kw)
    assert len(got) == len(ref)
    for a, b in zip(got, ref):
        assert a[0] == b[0] and a[3] == b[3] and a[4] == b[4] and a[5] == b[5]  # utts, mel lens, durations, texts: exact
        for ra, rb in zip(a[1], b[1]):
            assert abs(float(ra) - float(rb)) <= 2e-6 * float(rb)
        assert a[2].shape == b[2].shape
        assert (a[2].cpu() - b[2]).abs().max().item() < 3e-4  # log-mel, same bar as the MelSpec kernel test


def test_infer_batch_process_vs_oracle_and_batch1(cuda_dev):
    """All chunks in ONE padded solve == the reference's chunk-by-chunk batch-1 loop (isolate_samples)."""
    from f5e_tts_b200 import pipeline as P
    from f5e_tts_b200.model import CFM, DiT
    from f5e_tts_b200.vocos import Vocos
    cfg = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=256, conv_layers=2, text_num_embeds=256)
    sd = {k: v.to(torch.bfloat16).float() for k, v in O.make_state_dict(cfg, 31).items()}
    dit = DiT(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, mel_dim=100, text_num_embeds=256,
              text_dim=256, conv_layers=2)
    dit.load_state_dict(sd, strict=False)
    model = CFM(dit).to(cuda_dev).to(torch.bfloat16)  # vocab_char_map None: UTF-8 bytes are the tokens (list_str_to_tensor)
    voc = Vocos(seed=4321).to(cuda_dev)
    g = torch.Generator().manual_seed(2)
    sr = 16000
    ref_audio = 0.02 * torch.randn(2, int(1.1 * sr), generator=g)  # stereo, quiet (gain path), needs resampling
    ref_text = "short reference."
    chunks = ["first chunk of words, a bit longer than the others.", "tiny.", "and a third chunk to cross-fade into."]
    kw = dict(nfe_step=4, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=11)
    wave, out_sr, spec = next(P.infer_batch_process((ref_audio, sr), ref_text, chunks, model, voc, **kw))
    assert out_sr == 24000 and wave.dtype == np.float64
    # (a) the product itself, chunk by chunk at batch 1 (no padding anywhere): the batched solve must agree
    singles = [next(P.infer_batch_process((ref_audio, sr), ref_text, [c], model, voc, **kw)) for c in chunks]
    mel_cat = np.concatenate([s[2] for s in singles], axis=1)
    assert spec.shape == mel_cat.shape
    e_iso = np.linalg.norm(spec - mel_cat) / np.linalg.norm(mel_cat)
    print(f"batched vs batch-1 (product): mel rel-L2 {e_iso:.3e}")
    assert e_iso < 2e-3
    # (b) the oracle's restatement of the reference loop, fp32 and bf16
    rt = ref_text + " "
    ids = lambda s: list(s.encode("utf-8"))  # noqa: E731
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    vsd = voc.state_dict()  # the oracle decodes with the product vocoder's weights
    args = (vsd, ref_audio.to(cuda_dev), sr, ids(rt), [ids(c) for c in chunks], [len(ids(c)) for c in chunks], len(ids(rt)))
    w32, m32 = O.long_form(o32, *args, steps=4, seed=11)
    w16, m16 = O.long_form(o16, *args, steps=4, seed=11)
    assert wave.shape == w32.shape  # durations, chunk lengths and cross-fade bookkeeping: exact
    m32c, m16c = np.concatenate(m32, axis=1), np.concatenate([m.astype(np.float32) for m in m16], axis=1)
    e_got = np.linalg.norm(spec - m32c) / np.linalg.norm(m32c)
    e_ref = np.linalg.norm(m16c - m32c) / np.linalg.norm(m32c)
    print(f"long-form mel: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
    assert e_got <= 2.0 * e_ref + 5e-3
    w_got, w_ref = np.linalg.norm(wave - w32) / np.linalg.norm(w32), np.linalg.norm(w16 - w32) / np.linalg.norm(w32)
    print(f"long-form wave (gain + cross-fade included): product-vs-fp32 {w_got:.3e}  oracle-bf16-vs-fp32 {w_ref:.3e}")
    assert w_got <= 2.0 * w_ref + 2e-2


def test_eval_loop_vs_oracle(cuda_dev):
    """evaluate.infer_batches (batched decode, device-side gain, rank sharding) against the reference loop's restatement."""
    from f5e_tts_b200 import evaluate as E
    from f5e_tts_b200 import pipeline as P
    from f5e_tts_b200.model import CFM, DiT
    from f5e_tts_b200.vocos import Vocos
    cfg = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=256, conv_layers=2, text_num_embeds=256)
    sd = {k: v.to(torch.bfloat16).float() for k, v in O.make_state_dict(cfg, 41).items()}
    dit = DiT(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, mel_dim=100, text_num_embeds=256,
              text_dim=256, conv_layers=2)
    dit.load_state_dict(sd, strict=False)
    model = CFM(dit).to(cuda_dev).to(torch.bfloat16)
    voc = Vocos(seed=4321).to(cuda_dev)
    meta = _metainfo(7, 9)
    prompts = P.get_inference_prompt(meta, device=cuda_dev, infer_batch_size=1500, num_buckets=10, min_secs=1, max_secs=40)
    kw = dict(nfe_step=4, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=5)
    # two "ranks" in one process: together they cover every utterance exactly once
    got = dict(E.infer_batches(model, voc, prompts, rank=0, world=2, **kw) + E.infer_batches(model, voc, prompts, rank=1, world=2, **kw))
    assert sorted(got) == sorted(m[0] for m in meta)
    # keep_on_device (no per-batch D2H / host sync: what bench.py's eval-set leg gathers over NCCL): same bits, on the GPU
    on_dev = dict(E.infer_batches(model, voc, prompts, keep_on_device=True, **kw))
    for utt, w in got.items():
        assert on_dev[utt].is_cuda and w.device.type == "cpu" and torch.equal(on_dev[utt].cpu(), w), utt
    # oracle: same prompts, texts as UTF-8 byte ids (vocab_char_map None), fp32 and bf16
    oprompts = [(u, [float(r) for r in rms], mels.cpu(), rl, tl, [list(t.encode("utf-8")) for t in txt])
                for u, rms, mels, rl, tl, txt in prompts]
    vsd = voc.state_dict()
    r32 = O.eval_infer(O.OracleDiT(cfg, sd, torch.float32, cuda_dev), vsd, oprompts, steps=4, seed=5)
    r16 = O.eval_infer(O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev), vsd, oprompts, steps=4, seed=5)
    worst = []
    for utt, w in got.items():
        assert w.shape == r32[utt].shape and w.dtype == torch.float32  # lengths: exact
        e_got, e_ref = rel_l2(w, r32[utt]), rel_l2(r16[utt], r32[utt])
        print(f"{utt}: wave product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
        worst.append((e_got, e_ref, utt))
    for e_got, e_ref, utt in worst:
        assert e_got <= 2.0 * e_ref + 2e-2, utt
