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


# ---- Conformer encoder + content head -------------------------------------------------------------------------------
K_TOL = 2.0  # same bar as the DiT: err(product vs fp32 oracle) <= K_TOL * err(bf16 oracle vs fp32 oracle) + 5e-3 (rel-L2)


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def make_model(cfg, sd, dev):
    from f5e_tts_b200.ppg import ASRModel
    sd16 = {k: (v.to(torch.bfloat16).float() if v.dim() >= 2 and "global_cmvn" not in k else v) for k, v in sd.items()}
    m = ASRModel(cfg.input_dim, output_size=cfg.dim, attention_heads=cfg.heads, linear_units=cfg.ffn, num_blocks=cfg.blocks,
                 cnn_module_kernel=cfg.kernel, causal=cfg.causal, cnn_module_norm=cfg.cnn_norm)
    return m.load_state_dict(sd16).to(dev), sd16


@pytest.mark.parametrize("fixture", ["ppg_conformer", "ppg_conformer_causal_ln"])
def test_conformer_vs_reference_golden(cuda_dev, fixture):
    """The product against the output of the reference's own wenet modules (tests/golden/ppg_conformer*.pt): the default conv
    module (symmetric, BatchNorm) and the streaming recipe's (causal, LayerNorm)."""
    from f5e_tts_b200.ppg import PPGModelWapper
    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", fixture + ".pt"))
    cfg = PO.ConformerConfig(blocks=fx["blocks"], causal=fx.get("causal", False), cnn_norm=fx.get("cnn_norm", "batch_norm"))
    sd = PO.make_conformer_state_dict(cfg, seed=fx["seed"])
    model, sd16 = make_model(cfg, sd, cuda_dev)
    wrap = PPGModelWapper(model=model, device=cuda_dev)
    ppg, true_len = wrap.audio_to_ppg(fx["wave"].to(cuda_dev), sr=16000)
    assert ppg.shape == (1,) + tuple(fx["ppg"].shape) and true_len.tolist() == [fx["ppg"].shape[0]]
    o16 = PO.conformer_extract(cfg, sd16, fx["feats"].to(cuda_dev), torch.bfloat16).cpu()
    o32 = PO.conformer_extract(cfg, sd16, fx["feats"].to(cuda_dev)).cpu()
    e_got, e_ref, e_w = rel_l2(ppg[0].cpu(), o32), rel_l2(o16, o32), rel_l2(o32, fx["ppg"])
    print(f"golden: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}  (bf16 weights vs reference fp32 weights {e_w:.3e})")
    assert e_got <= K_TOL * e_ref + 5e-3
    assert rel_l2(ppg[0].cpu(), fx["ppg"]) <= K_TOL * rel_l2(o16, fx["ppg"]) + 5e-3


@pytest.mark.parametrize("blocks", [1, 12])
def test_conformer_ragged_batch_vs_oracle(cuda_dev, blocks):
    from f5e_tts_b200.ppg import PPGModelWapper
    cfg = PO.ConformerConfig(blocks=blocks)
    sd = PO.make_conformer_state_dict(cfg, seed=5)
    model, sd16 = make_model(cfg, sd, cuda_dev)
    wrap = PPGModelWapper(model=model, device=cuda_dev)
    g = torch.Generator().manual_seed(3)
    lens = torch.tensor([48000, 17000, 33333, 1000])
    wave = 0.1 * torch.randn(4, 48000, generator=g)
    ppg, true_len = wrap.audio_to_ppg(wave.to(cuda_dev), sr=16000, lens=lens.to(cuda_dev))
    worst = 0.0
    for b in range(4):
        w = wave[b, :int(lens[b])]
        want32, n = PO.audio_to_ppg(cfg, sd16, w.to(cuda_dev))
        want16, _ = PO.audio_to_ppg(cfg, sd16, w.to(cuda_dev), torch.bfloat16)
        assert int(true_len[b]) == n
        assert float(ppg[b, n:].abs().max()) == 0.0 if n < ppg.shape[1] else True
        e_got, e_ref = rel_l2(ppg[b, :n], want32), rel_l2(want16, want32)
        print(f"blocks={blocks} utt {b} ({n} frames): product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
        assert e_got <= K_TOL * e_ref + 5e-3
        worst = max(worst, e_got / (K_TOL * e_ref + 5e-3))
        alone, n1 = wrap.audio_to_ppg(w.to(cuda_dev)[None], sr=16000)  # ragged batch == batch-1 call, bit for bit
        assert int(n1[0]) == n and torch.equal(alone[0, :n], ppg[b, :n])
    print(f"blocks={blocks}: worst product error / allowed = {worst:.2f}")
    # the CUDA-graph path replays the same kernels: same bits, also on new data of the same shape
    wave2, lens2 = wave.flip(0).contiguous(), torch.tensor([20000, 48000, 999, 31000])
    want2, _ = wrap.audio_to_ppg(wave2.to(cuda_dev), sr=16000, lens=lens2.to(cuda_dev))
    model.use_cuda_graph = True
    g1, n1 = wrap.audio_to_ppg(wave.to(cuda_dev), sr=16000, lens=lens.to(cuda_dev))
    g2, _ = wrap.audio_to_ppg(wave2.to(cuda_dev), sr=16000, lens=lens2.to(cuda_dev))
    assert torch.equal(g1, ppg) and torch.equal(n1, true_len) and torch.equal(g2, want2)


def test_kernels_of_the_conformer(cuda_dev):
    """subsample conv, rel-pos key bias + attention, GLU/depthwise conv: each against plain torch on random data."""
    import math
    import torch.nn.functional as F
    from f5e_tts_b200 import _lib
    from f5e_tts_b200._lib import check, ptr, stream_ptr
    lib, C = _lib.load(), _lib.C
    g = torch.Generator().manual_seed(1)
    # 1. Conv2d(1, C, 3, 2) + ReLU
    B, T, Fd, Cc = 2, 21, 80, 256
    feats = torch.randn(B, T, Fd, generator=g).to(cuda_dev)
    w, b = torch.randn(Cc, 1, 3, 3, generator=g).to(cuda_dev) / 3, torch.randn(Cc, generator=g).to(cuda_dev)
    nfr = torch.tensor([21, 10], dtype=torch.int32, device=cuda_dev)
    t_pad, Fo = 16, 39
    out = torch.empty(B * t_pad, Fo * Cc, device=cuda_dev, dtype=torch.bfloat16)
    check(lib.f5e_subsample_conv(ptr(feats), C.c_int64(feats.stride(0)), ptr(nfr), B, T, Fd, ptr(w.view(Cc, 9).contiguous()), ptr(b), Cc,
                                 ptr(out), t_pad, stream_ptr()), "f5e_subsample_conv")
    want = F.relu(F.conv2d(feats[:, None], w, b, stride=2)).permute(0, 2, 3, 1)  # [B, t, f, c]
    got = out.view(B, t_pad, Fo, Cc).float()
    assert torch.equal(got[0, :10], want[0].bfloat16().float()) or (got[0, :10] - want[0]).abs().max() < 2e-2
    assert (got[1, :4] - want[1, :4]).abs().max() < 2e-2 and float(got[1, 4:].abs().max()) == 0.0 and float(got[0, 10:].abs().max()) == 0.0
    # 2. rel-pos attention = q.(k+p) + (u.k + v.p)
    S, n_pad, H = 2, 200, 4
    d = H * 64
    qkv = (torch.randn(S * n_pad, 3 * d, generator=g) * 0.5).to(cuda_dev).bfloat16()
    pos = (torch.randn(n_pad, d, generator=g) * 0.5).to(cuda_dev)
    u, v = (torch.randn(H, 64, generator=g) * 0.3).to(cuda_dev), (torch.randn(H, 64, generator=g) * 0.3).to(cuda_dev)
    klen = torch.tensor([200, 77], dtype=torch.int32, device=cuda_dev)
    q, k, vv = [t.float().view(S, n_pad, H, 64).transpose(1, 2) for t in qkv.split(d, dim=1)]
    p = pos.view(n_pad, H, 64).transpose(0, 1)
    scores = ((q + u[None, :, None]) @ k.transpose(-1, -2) + (q + v[None, :, None]) @ p.transpose(-1, -2)[None]) / 8.0
    scores = scores.masked_fill(torch.arange(n_pad, device=cuda_dev)[None, None, None] >= klen[:, None, None, None], float("-inf"))
    want = (torch.softmax(scores, -1) @ vv).transpose(1, 2).reshape(S, n_pad, d)
    ld_bias = 256
    kb = torch.zeros(S * H, ld_bias, device=cuda_dev)
    work = qkv.clone()
    check(lib.f5e_relpos_prepare(ptr(work), work.stride(0), ptr(pos), ptr(u), ptr(v), ptr(kb), ld_bias, S, n_pad, H, stream_ptr()),
          "f5e_relpos_prepare")
    att = torch.zeros(S * n_pad, d, device=cuda_dev, dtype=torch.bfloat16)
    check(lib.f5e_attention_bias(ptr(work), work.stride(0), ptr(att), att.stride(0), ptr(klen), ptr(kb), ld_bias, S, n_pad, H,
                                 C.c_float(0.125), stream_ptr()), "f5e_attention_bias")
    got = att.view(S, n_pad, d).float()
    for s_ in range(S):
        n = int(klen[s_])
        e = rel_l2(got[s_, :n], want[s_, :n])
        print(f"rel-pos attention sample {s_}: rel-L2 {e:.3e}")
        assert e < 2e-2
    # 3. GLU + depthwise conv + swish
    S, n_pad, Cc, taps = 2, 50, 256, 15
    gin = torch.randn(S * n_pad, 2 * Cc, generator=g).to(cuda_dev).bfloat16()
    wd, bd = (torch.randn(Cc, taps, generator=g) / 4).to(cuda_dev), torch.randn(Cc, generator=g).to(cuda_dev)
    ln_ = torch.tensor([50, 23], dtype=torch.int32, device=cuda_dev)
    out = torch.empty(S * n_pad, Cc, device=cuda_dev, dtype=torch.bfloat16)
    for lpad in (7, 14):
        check(lib.f5e_glu_dwconv(ptr(gin), ptr(wd), ptr(bd), ptr(ln_), ptr(out), S, n_pad, Cc, taps, lpad, None, None, None,
                                 C.c_float(0.0), stream_ptr()), "f5e_glu_dwconv")
        for s_ in range(S):
            n = int(ln_[s_])
            x = F.glu(gin.view(S, n_pad, 2 * Cc)[s_, :n].float(), dim=-1).t()[None]
            y = F.conv1d(F.pad(x, (lpad, taps - 1 - lpad)), wd[:, None], bd, groups=Cc)
            y = (y * torch.sigmoid(y))[0].t()
            got = out.view(S, n_pad, Cc)[s_].float()
            assert rel_l2(got[:n], y) < 1e-2 and (float(got[n:].abs().max()) == 0.0 if n < n_pad else True)
    # causal left context = a per-channel constant, LayerNorm over channels before the swish
    pv, lw, lb = torch.randn(Cc, generator=g).to(cuda_dev), (1 + 0.1 * torch.randn(Cc, generator=g)).to(cuda_dev), torch.randn(Cc, generator=g).to(cuda_dev)
    check(lib.f5e_glu_dwconv(ptr(gin), ptr(wd), ptr(bd), ptr(ln_), ptr(out), S, n_pad, Cc, taps, taps - 1, ptr(pv), ptr(lw), ptr(lb),
                             C.c_float(1e-5), stream_ptr()), "f5e_glu_dwconv")
    for s_ in range(S):
        n = int(ln_[s_])
        x = F.glu(gin.view(S, n_pad, 2 * Cc)[s_, :n].float(), dim=-1).t()
        x = torch.cat([pv[:, None].expand(Cc, taps - 1), x], dim=1)[None]
        y = F.layer_norm(F.conv1d(x, wd[:, None], bd, groups=Cc)[0].t(), (Cc,), lw, lb, 1e-5)
        y = y * torch.sigmoid(y)
        assert rel_l2(out.view(S, n_pad, Cc)[s_, :n].float(), y) < 1e-2


def test_voice_conversion_end_to_end(cuda_dev):
    """Source audio -> kaldi fbank -> Conformer PPG -> CFM.sample_vc (eval_infer_batch_vc.py:215-225), product chain against the
    oracle chain under the bf16 bar."""
    import f5e_oracle as O
    from test_parity_gpu import build
    from f5e_tts_b200.ppg import PPGModelWapper
    cfg = O.DiTConfig(dim=512, depth=2, heads=8, text_dim=256, conv_layers=2, text_num_embeds=50, text_mask_padding=False,
                      pe_attn_head=1, use_ppg=True, ppg_dim=256)
    model, sd = build(cfg, 41, cuda_dev)
    pcfg = PO.ConformerConfig(blocks=2)
    psd = PO.make_conformer_state_dict(pcfg, seed=6)
    pmodel, psd16 = make_model(pcfg, psd, cuda_dev)
    wrap = PPGModelWapper(model=pmodel, device=cuda_dev)
    g = torch.Generator().manual_seed(42)
    B = 2
    src_len = torch.tensor([40000, 29000])
    src = 0.1 * torch.randn(B, 40000, generator=g)
    ref_len = torch.tensor([60, 45], device=cuda_dev)
    dur = torch.tensor([290, 230], device=cuda_dev)
    cond = torch.randn(B, 60, cfg.mel_dim, generator=g).to(cuda_dev)
    ppg, true_len = wrap.audio_to_ppg(src.to(cuda_dev), sr=16000, lens=src_len.to(cuda_dev))
    o_ppg = {}
    for dt in (torch.float32, torch.bfloat16):
        rows = [PO.audio_to_ppg(pcfg, psd16, src[b, :int(src_len[b])].to(cuda_dev), dt)[0] for b in range(B)]
        o_ppg[dt] = torch.nn.utils.rnn.pad_sequence(rows, batch_first=True)
    assert o_ppg[torch.float32].shape == ppg.shape
    kw = dict(steps=6, sway_sampling_coef=-1.0)
    okw = dict(mode="vc", alpha_spk=2.5, alpha_cont=3.0, **kw)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    out16, traj16, aux = O.cfm_sample(o16, cond, None, dur, ppg=o_ppg[torch.bfloat16], lens=ref_len, seed=3, return_aux=True, **okw)
    out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), None, dur, ppg=o_ppg[torch.float32], lens=ref_len, y0=aux["y0"].float(), **okw)
    got, traj = model.sample_vc(cond, ppg, dur, lens=ref_len, seed=3, alpha_spk=2.5, alpha_ppg=3.0, **kw)
    for i in range(B):
        L = int(aux["duration"][i])
        e_got, e_ref = rel_l2(traj[-1][i, :L], traj32[-1][i, :L]), rel_l2(traj16[-1][i, :L], traj32[-1][i, :L])
        print(f"vc end to end sample {i}: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
        assert e_got <= K_TOL * e_ref + 5e-3


def test_get_vc_inference_prompt_and_eval_loop(cuda_dev):
    """pipeline.get_vc_inference_prompt (eval/utils_eval.py:245-397) against the oracle's one-utterance-at-a-time restatement,
    then the sample_vc evaluation loop (eval_infer_batch_vc.py:203-232) over its batches."""
    import f5e_oracle as O
    from test_parity_gpu import build
    from f5e_tts_b200.evaluate import infer_batches
    from f5e_tts_b200.pipeline import get_vc_inference_prompt
    from f5e_tts_b200.ppg import PPGModelWapper
    from f5e_tts_b200.vocos import Vocos
    pcfg = PO.ConformerConfig(blocks=2)
    psd = PO.make_conformer_state_dict(pcfg, seed=8)
    pmodel, psd16 = make_model(pcfg, psd, cuda_dev)
    wrap = PPGModelWapper(model=pmodel, device=cuda_dev)
    g = torch.Generator().manual_seed(77)
    meta = []
    for i, (psr, gsr, ch) in enumerate([(24000, 16000, 1), (16000, 24000, 1), (24000, 16000, 2), (24000, 24000, 1), (16000, 16000, 1)]):
        pw = 0.05 * torch.randn(1, int(psr * (3.0 + 0.3 * i)), generator=g)
        gw = 0.1 * torch.randn(ch, int(gsr * (2.0 + 0.5 * i)), generator=g)
        meta.append((f"utt{i}", "some prompt text here.", (pw, psr), "words to say " * (2 + i), (gw, gsr)))
    got = get_vc_inference_prompt(meta, wrap, infer_batch_size=1500, device=cuda_dev)
    want32 = PO.get_vc_inference_prompt(meta, pcfg, psd16, infer_batch_size=1500)
    want16 = PO.get_vc_inference_prompt(meta, pcfg, psd16, infer_batch_size=1500, dtype=torch.bfloat16)
    assert len(got) == len(want32) and sum(len(b[0]) for b in got) == len(meta)
    for gb, w32, w16 in zip(got, want32, want16):
        assert gb[0] == w32[0] and gb[3] == w32[3] and gb[4] == w32[4]  # utterances, ref_mel_lens, total_mel_lens: exact
        assert gb[2].shape == w32[2].shape and (gb[2].cpu() - w32[2]).abs().max() < 3e-4
        assert gb[5].shape == w32[5].shape
        for r in range(gb[5].shape[0]):
            e_got, e_ref = rel_l2(gb[5][r].cpu(), w32[5][r]), rel_l2(w16[5][r], w32[5][r])
            print(f"{gb[0][r]}: ppg {tuple(gb[5][r].shape)} product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
            assert e_got <= K_TOL * e_ref + 5e-3
    cfg = O.DiTConfig(dim=512, depth=2, heads=8, text_dim=256, conv_layers=2, text_num_embeds=50, text_mask_padding=False,
                      pe_attn_head=1, use_ppg=True, ppg_dim=256)
    model, _ = build(cfg, 43, cuda_dev)
    voc = Vocos(seed=4321).to(cuda_dev)
    kw = dict(nfe_step=4, sway_sampling_coef=-1.0, seed=5, sampler="sample_vc", sampler_kwargs=dict(alpha_spk=2.5, alpha_ppg=3.0))
    out = infer_batches(model, voc, got, device=cuda_dev, **kw)
    assert sorted(u for u, _ in out) == sorted(m[0] for m in meta)
    by_utt = dict(out)
    # the loop's waves are what a direct sample_vc + decode of the same batch gives
    utts, rms, mels, ref_lens, tot_lens, ppg = got[0]
    rl, tl = torch.tensor(ref_lens, device=cuda_dev), torch.tensor(tot_lens, device=cuda_dev)
    gen, _ = model.sample_vc(cond=mels.to(cuda_dev), ppg=ppg, duration=tl, lens=rl, steps=4, sway_sampling_coef=-1.0, seed=5,
                             alpha_spk=2.5, alpha_ppg=3.0)
    wav = voc.decode_varlen(gen, rl, tl, max_frames=max(t - r for t, r in zip(tot_lens, ref_lens)))
    n0 = (tot_lens[0] - ref_lens[0] - 1) * 256
    w0 = wav[0, :n0].cpu()
    if float(rms[0]) < 0.1:
        w0 = w0 * float(rms[0]) / 0.1
    assert by_utt[utts[0]].shape == (1, n0) and torch.allclose(by_utt[utts[0]][0], w0, atol=1e-6)
    two = infer_batches(model, voc, got, device=cuda_dev, rank=0, world=2, **kw) + infer_batches(model, voc, got, device=cuda_dev, rank=1, world=2, **kw)
    assert sorted(u for u, _ in two) == sorted(by_utt) and all(torch.equal(w, by_utt[u]) for u, w in two)


def test_ppg_map_output(cuda_dev):
    """PPGModelWapper(output_type="map") with the tables passed as arguments: softmax(ppg @ w^T + b) @ global_phn_center on
    the fp32-accurate GEMMs + the row-softmax kernel, against the reference-generated fixture (and the oracle at a wider
    shape); padded frames exactly zero."""
    import os

    from f5e_tts_b200.ppg import PPGModelWapper
    from oracle import ppg_oracle as PO

    d = torch.load(os.path.join(os.path.dirname(__file__), "golden", "ppg_map.pt"), weights_only=False)
    for ratio, key in ((1.0, "out_1"), (0.35, "out_035")):
        wr = object.__new__(PPGModelWapper)   # the table handling of __init__ without an ASR checkpoint
        PPGModelWapper.__init__(wr, model=_tiny_asr(cuda_dev), device=cuda_dev, output_type="map", map_mix_ratio=ratio,
                                global_phn_center=d["center"], para_softmax=dict(w=d["w"], b=d["b"]))
        got = wr.ppg_to_target(d["ppg"].to(cuda_dev), d["true_len"].to(cuda_dev))
        assert torch.allclose(got.cpu(), d[key], atol=2e-5, rtol=1e-4), (got.cpu() - d[key]).abs().max()
        assert (got[1, 17:] == 0).all()
    g = torch.Generator().manual_seed(5)
    B, T, dim, V = 3, 470, 256, 601
    ppg = torch.randn(B, T, dim, generator=g)
    tl = torch.tensor([470, 300, 1])
    w, b, c = torch.randn(V, dim, generator=g) / 16, torch.randn(V, generator=g), torch.randn(V, dim, generator=g)
    wr = object.__new__(PPGModelWapper)
    PPGModelWapper.__init__(wr, model=_tiny_asr(cuda_dev), device=cuda_dev, output_type="map", global_phn_center=c,
                            para_softmax=dict(w=w, b=b))
    got = wr.ppg_to_target(ppg.to(cuda_dev), tl.to(cuda_dev)).cpu()
    ref = PO.ppg_to_target(ppg.double(), tl, "map", w.double(), b.double(), c.double(), 1.0)
    assert (got - ref).abs().max() < 5e-5
    with pytest.raises(ValueError):
        PPGModelWapper(model=_tiny_asr(cuda_dev), device=cuda_dev, output_type="map")


def _tiny_asr(dev):
    from f5e_tts_b200.ppg import ASRModel
    return ASRModel(80, output_size=256, attention_heads=4, linear_units=256, num_blocks=1)
