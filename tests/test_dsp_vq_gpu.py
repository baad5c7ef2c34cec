"""GPU parity of the fp32 front / back ends and the codebook call, through the C ABI, against the oracle."""
import os

import pytest
import torch

from oracle import f5e_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


# ---- MelSpec: tolerance 2e-4 absolute on the log-mel (fp32 FFT vs torch.stft; values are O(1..5)) ----------------------
MEL_ATOL = 2e-4


@pytest.mark.parametrize("B,nw", [(1, 24000), (3, 83968), (2, 5000), (1, 513), (2, 256 * 40 + 17)])
def test_melspec_vs_oracle(cuda_dev, B, nw):
    from f5e_tts_b200.frontend import MelSpec
    g = torch.Generator().manual_seed(nw)
    wav = 0.1 * torch.randn(B, nw, generator=g)
    got = MelSpec()(wav.to(cuda_dev))
    ref = O.mel_spectrogram(wav)
    assert got.shape == ref.shape == (B, 100, nw // 256 + 1)
    assert (got.cpu() - ref).abs().max().item() < MEL_ATOL, (got.cpu() - ref).abs().max()


def test_melspec_golden_fixture(cuda_dev):
    from f5e_tts_b200.frontend import MelSpec
    d = torch.load(os.path.join(G, "melspec.pt"), weights_only=False)  # output of the reference's MelSpec itself
    got = MelSpec()(d["wav"].to(cuda_dev)).cpu()
    assert (got - d["mel"]).abs().max().item() < MEL_ATOL
    # [B,1,nw] input form (modules.py:134-135)
    got3 = MelSpec()(d["wav"].unsqueeze(1).to(cuda_dev)).cpu()
    assert torch.equal(got3, got)


def test_melspec_silence_is_clamped(cuda_dev):
    from f5e_tts_b200.frontend import MelSpec
    got = MelSpec()(torch.zeros(1, 4096, device=cuda_dev))
    assert torch.allclose(got, torch.full_like(got, float(torch.log(torch.tensor(1e-5)))))


# ---- ISTFT head vs torch.istft -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,T", [(1, 50), (2, 611), (3, 9), (1, 2)])
def test_istft_head_vs_torch(cuda_dev, B, T):
    import ctypes as C
    from f5e_tts_b200 import _lib
    from f5e_tts_b200.dsp import fft_tables
    g = torch.Generator().manual_seed(T)
    x = torch.randn(B, 1026, T, generator=g)
    x[:, 513:] *= 3.0          # phases of several radians
    x[:, :513] *= 2.0          # some magnitudes hit the 1e2 clip
    ref = O.istft_head(x)
    spec = x.permute(0, 2, 1).contiguous().to(cuda_dev)  # [B, T, 1026]
    window, tw = fft_tables(cuda_dev)
    L = (T - 1) * 256
    out = torch.full((B, L + 7), 123.0, device=cuda_dev)
    _lib.check(_lib.load().f5e_istft_head(_lib.ptr(spec), C.c_int64(1026), B, T, None, _lib.ptr(window), _lib.ptr(tw), _lib.ptr(out),
                                          C.c_int64(out.stride(0)), L, _lib.stream_ptr()), "f5e_istft_head")
    assert torch.all(out[:, L:] == 123.0)  # nothing written past out_len
    err = (out[:, :L].cpu() - ref).abs().max().item()
    assert err < 2e-4 * max(1.0, ref.abs().max().item()), err
    assert rel_l2(out[:, :L].cpu(), ref) < 1e-5


# ---- Vocos decode (fp32 activations, bf16x3 GEMMs): rel-L2 <= 1e-4, max-abs <= 1e-3 * peak ---------------------------
def _vocos(cuda_dev, seed=4321):
    from f5e_tts_b200.vocos import Vocos
    sd = O.make_vocos_state_dict(seed)
    v = Vocos()
    v.load_state_dict(sd)
    return v.to(cuda_dev), sd


@pytest.mark.parametrize("B,T", [(1, 100), (2, 611), (4, 37)])
def test_vocos_decode_vs_oracle(cuda_dev, B, T):
    voc, sd = _vocos(cuda_dev)
    g = torch.Generator().manual_seed(B * 1000 + T)
    mel = torch.randn(B, 100, T, generator=g) * 2 - 3
    ref = O.vocos_decode(sd, mel)
    got = voc.decode(mel.to(cuda_dev)).cpu()
    assert got.shape == ref.shape == (B, (T - 1) * 256)
    assert rel_l2(got, ref) < 1e-4, rel_l2(got, ref)
    assert (got - ref).abs().max().item() < 1e-3 * ref.abs().max().item()


def test_vocos_decode_varlen_equals_per_utterance(cuda_dev):
    """The batched varlen decode reproduces the reference's one-utterance-at-a-time loop (eval_infer_batch.py:208-212)."""
    voc, sd = _vocos(cuda_dev)
    g = torch.Generator().manual_seed(5)
    B, N = 3, 300
    out = (torch.randn(B, N, 100, generator=g) * 2 - 3).to(cuda_dev)  # CFM.sample layout [B, N, 100]
    start = torch.tensor([40, 0, 120], device=cuda_dev)
    end = torch.tensor([300, 17, 251], device=cuda_dev)
    wav = voc.decode_varlen(out.bfloat16(), start, end)
    assert wav.shape == (B, (260 - 1) * 256)
    for i in range(B):
        m = out.bfloat16()[i, int(start[i]): int(end[i])].permute(1, 0).float().unsqueeze(0)
        ref = O.vocos_decode(sd, m.cpu())[0]
        L = ref.numel()
        assert rel_l2(wav[i, :L].cpu(), ref) < 1e-4, (i, rel_l2(wav[i, :L].cpu(), ref))
        assert wav[i, L:].abs().max().item() == 0 if L < wav.shape[1] else True


# ---- VQ: indices bit exact --------------------------------------------------------------------------------------------
def _quantizer(cuda_dev, dim, V, Gr, seed):
    from f5e_tts_b200.quantizer import GumbelVectorQuantizer
    torch.manual_seed(seed)
    q = GumbelVectorQuantizer(dim=dim, num_vars=V, temp=(2.0, 0.5, 0.999), groups=Gr, combine_groups=False, vq_dim=dim,
                              time_first=True).eval()
    sd = {"quantizer.vars": q.vars.data.clone(), "quantizer.weight_proj.weight": q.weight_proj.weight.data.clone(),
          "quantizer.weight_proj.bias": q.weight_proj.bias.data.clone()}
    return q.to(cuda_dev), sd


def test_vq_golden_fixture_bit_exact(cuda_dev):
    from f5e_tts_b200.quantizer import GumbelVectorQuantizer
    d = torch.load(os.path.join(G, "vq.pt"), weights_only=False)  # outputs of the reference's quantizer itself
    q = GumbelVectorQuantizer(dim=64, num_vars=10, temp=(2.0, 0.5, 0.999), groups=2, combine_groups=False, vq_dim=64,
                              time_first=True).eval()
    q.load_state_dict({k.replace("quantizer.", ""): v for k, v in d["sd"].items()})
    r = q.to(cuda_dev)(d["x"].to(cuda_dev), produce_targets=True)
    assert torch.equal(r["targets"].cpu(), d["targets"])
    assert torch.equal(r["x"].cpu(), d["xq"])
    assert abs(float(r["code_perplexity"]) - float(d["code_perplexity"])) < 1e-4
    assert abs(float(r["prob_perplexity"]) - float(d["prob_perplexity"])) < 1e-4


@pytest.mark.parametrize("B,T", [(2, 100), (32, 940)])
def test_vq_indices_bit_exact_vs_oracle(cuda_dev, B, T):
    q, sd = _quantizer(cuda_dev, 512, 100, 2, 9)
    g = torch.Generator().manual_seed(B + T)
    x = torch.randn(B, T, 512, generator=g)
    ref = O.vq_forward({k: v for k, v in sd.items()}, x, groups=2, num_vars=100)
    xq, tg = q.forward_idx(x.to(cuda_dev))
    assert tg.dtype == torch.int64 and torch.equal(tg.cpu(), ref["targets"])
    assert torch.equal(xq.cpu(), ref["x"])
    r = q(x.to(cuda_dev))
    assert "targets" not in r
    assert abs(float(r["code_perplexity"]) - float(ref["code_perplexity"])) < 1e-3 * float(ref["code_perplexity"])
    assert abs(float(r["prob_perplexity"]) - float(ref["prob_perplexity"])) < 1e-3 * float(ref["prob_perplexity"])


def test_vq_near_ties_are_redecided_in_fp32(cuda_dev):
    """Rows built so that two codes of a group tie to within 1e-6: the kernel must take the sequential-fp32 decision."""
    q, sd = _quantizer(cuda_dev, 512, 100, 2, 11)
    with torch.no_grad():
        q.weight_proj.weight[7] = q.weight_proj.weight[3]            # exact duplicates: first index must win
        q.weight_proj.weight[150] = q.weight_proj.weight[120] * (1 + 1e-7)
    q._plan = None
    w = q.weight_proj.weight.detach().cpu()
    x = torch.randn(4, 64, 512)
    x[0, :, :] = w[3] * 0.05 + 0.01 * torch.randn(64, 512)          # makes codes 3/7 the winners of group 0
    x[1, :, :] = w[120] * 0.05 + 0.01 * torch.randn(64, 512)
    xq, tg = q.forward_idx(x.to(cuda_dev))
    logits = torch.nn.functional.linear(x.reshape(-1, 512).double(), w.double()).view(-1, 2, 100)
    ref = logits.argmax(-1).view(4, 64, 2)  # fp64 decision; duplicates -> first index
    assert int(q.last_refined) > 0
    assert torch.equal(tg.cpu()[0, :, 0], torch.full((64,), 3))
    mism = (tg.cpu() != ref)
    # every disagreement with the fp64 decision must be a genuine sub-fp32 tie
    top2 = logits.view(4, 64, 2, 100).topk(2, -1).values
    margin = (top2[..., 0] - top2[..., 1]) / top2[..., 0].abs().clamp_min(1)
    assert bool((margin[mism] < 1e-5).all())


@pytest.mark.parametrize("B,T,groups,num_vars", [(32, 940, 2, 100), (3, 77, 1, 128), (2, 130, 2, 7)])
def test_vq_fused_decision_equals_unfused(cuda_dev, B, T, groups, num_vars):
    """north_star (3): arg-max + gather fused into the logit GEMM's epilogue (f5e_vq_quantize_fused, no logits round trip)
    gives the SAME indices and code vectors as the unfused path (GEMM -> fp32 logits -> vq_select), bit for bit, incl. rows
    sent through the sequential-fp32 re-decision; the perplexities agree to fp32 summation order."""
    q, sd = _quantizer(cuda_dev, 512, num_vars, groups, 21)
    with torch.no_grad():
        if num_vars > 5:
            q.weight_proj.weight[4] = q.weight_proj.weight[1]      # exact duplicate rows: ties, first index must win
    q._plan = None
    g = torch.Generator().manual_seed(B * T)
    x = torch.randn(B, T, 512, generator=g).to(cuda_dev)
    x[0, :8] = q.weight_proj.weight[1].detach() * 0.05
    assert q.fused_decision and q._get_plan()["fused"] is not None
    r1 = q(x, produce_targets=True)
    n1 = int(q.last_refined)
    q.fused_decision = False
    r0 = q(x, produce_targets=True)
    n0 = int(q.last_refined)
    assert torch.equal(r1["targets"], r0["targets"])
    assert torch.equal(r1["x"], r0["x"])
    assert n1 == n0 and (num_vars <= 5 or n1 >= 8)
    for k in ("code_perplexity", "prob_perplexity"):
        assert abs(float(r1[k]) - float(r0[k])) <= 1e-4 * abs(float(r0[k])), k
    # bf16 output path (the DiT feeds bf16 embeddings)
    q.fused_decision = True
    rb = q(x.bfloat16(), produce_targets=True)
    q.fused_decision = False
    rb0 = q(x.bfloat16(), produce_targets=True)
    assert rb["x"].dtype == torch.bfloat16 and torch.equal(rb["x"], rb0["x"]) and torch.equal(rb["targets"], rb0["targets"])
