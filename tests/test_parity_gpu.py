"""GPU parity of the product path (through the C ABI) against the oracle on the same seeded weights / inputs.

Bar (north_star): integer/bool side products bit exact; floating point within a stated bf16 tolerance.  The
tolerance is anchored on the oracle itself: the product (bf16 operands, fp32 accumulation, fused epilogues) must
be no further from the fp32 oracle than K_TOL x the distance of the SAME oracle run in bf16 (what the reference
does on a GPU in bf16), plus an absolute floor.
"""
import pytest
import torch

from oracle import f5e_oracle as O

pytestmark = pytest.mark.gpu
K_TOL = 2.0


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def build(cfg, seed, dev):
    from f5e_tts_b200.model import CFM, DiT
    sd = O.make_state_dict(cfg, seed)
    sd = {k: v.to(torch.bfloat16).float() for k, v in sd.items()}  # bf16-representable weights on both sides
    ppg_config = dict(use_ppg=cfg.use_ppg, ppg_dim=cfg.ppg_dim, use_transformer=cfg.ppg_transformer,
                      transformer_config=dict(nhead=cfg.ppg_nhead, dim_feedforward=cfg.ppg_ff, dropout=0.1, num_layers=cfg.ppg_layers))
    cb_config = dict(use_codebook=cfg.use_codebook, num_vars=cfg.cb_num_vars, temp_start=2.0, temp_stop=0.5,
                     temp_decay=0.999, groups=cfg.cb_groups, combine_groups=False, weight_proj_depth=1, weight_proj_factor=1)
    dit = DiT(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, mel_dim=cfg.mel_dim,
              text_num_embeds=cfg.text_num_embeds, text_dim=cfg.text_dim, text_mask_padding=cfg.text_mask_padding,
              conv_layers=cfg.conv_layers, pe_attn_head=cfg.pe_attn_head, qk_norm=cfg.qk_norm,
              long_skip_connection=cfg.long_skip_connection, ppg_config=ppg_config, cb_config=cb_config)
    missing, unexpected = dit.load_state_dict(sd, strict=False)
    assert not unexpected and all("num_batches_tracked" in k or "inv_freq" in k for k in missing), (missing, unexpected)
    model = CFM(dit, ppg_config=ppg_config, cb_config=cb_config).to(dev).to(torch.bfloat16)
    return model, sd


def inputs(cfg, B, N, seed, dev, ppg=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, N, cfg.mel_dim, generator=g)
    cond = torch.randn(B, N, cfg.mel_dim, generator=g)
    text = torch.randint(0, cfg.text_num_embeds, (B, 23), generator=g)
    if B > 1:
        text[1, 15:] = -1
    p = torch.randn(B, N // 2, cfg.ppg_dim, generator=g) if ppg else None
    return x.to(dev), cond.to(dev), text.to(dev), (p.to(dev) if ppg else None)


SMALL = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=256, conv_layers=2, text_num_embeds=50)
SMALL_PPG = O.DiTConfig(dim=768, depth=2, heads=12, text_dim=256, conv_layers=2, text_num_embeds=50, text_mask_padding=False,
                        pe_attn_head=1, use_ppg=True, ppg_dim=64)
BASE2 = O.DiTConfig(depth=2)  # F5TTS_v1_Base widths, 2 layers
# PPGEmbedding(use_transformer=True) (reference dit.py:105-119): nn.TransformerEncoder over the PPG stream
PPG_TR = O.DiTConfig(dim=512, depth=2, heads=8, text_dim=256, conv_layers=2, text_num_embeds=50, text_mask_padding=False,
                     pe_attn_head=1, use_ppg=True, ppg_dim=256, ppg_transformer=True, ppg_nhead=4, ppg_ff=512, ppg_layers=2)


# qk_norm="rms_norm" (modules.py:391-398, 463-467) + long_skip_connection (dit.py:264, 456-467) + partial-head RoPE
QKN_SKIP = O.DiTConfig(dim=512, depth=3, heads=8, text_dim=256, conv_layers=2, text_num_embeds=50, qk_norm="rms_norm",
                       long_skip_connection=True, pe_attn_head=3)


@pytest.mark.parametrize("cfg,B,N,lens", [(SMALL, 1, 200, None), (SMALL, 3, 333, [333, 150, 290]),
                                           (QKN_SKIP, 1, 200, None), (QKN_SKIP, 3, 333, [333, 150, 290]),
                                           (SMALL_PPG, 2, 260, [260, 131]), (BASE2, 2, 300, [300, 177]),
                                           (PPG_TR, 2, 260, [260, 131]),
                                           (BASE2, 2, 2812, [2812, 1903])])  # BASELINE configs[4] length (30 s utterances)
def test_dit_eval_vs_oracle(cuda_dev, cfg, B, N, lens):
    model, sd = build(cfg, 7, cuda_dev)
    x, cond, text, ppg = inputs(cfg, B, N, 8, cuda_dev, cfg.use_ppg)
    mask = None if lens is None else O.lens_to_mask(torch.tensor(lens, device=cuda_dev), N)
    t = torch.tensor(0.37, device=cuda_dev)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    for drop in (False, True):
        ref = o32.sample(x, cond, text, ppg, t, drop, drop, drop, mask)
        ref16 = o16.sample(x.bfloat16(), cond.bfloat16(), text, ppg if ppg is None else ppg.bfloat16(), t.bfloat16(), drop, drop, drop, mask)
        got = model.transformer.sample(x.bfloat16(), cond.bfloat16(), text, ppg, t.bfloat16(), drop, drop, drop, mask)
        model.transformer.clear_cache()
        for i in range(B):
            L = N if lens is None else lens[i]
            e_got, e_ref = rel_l2(got[i, :L], ref[i, :L]), rel_l2(ref16[i, :L], ref[i, :L])
            print(f"cfg dim={cfg.dim} B={B} drop={drop} sample {i}: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
            assert e_got <= K_TOL * e_ref + 5e-3, (i, e_got, e_ref)


@pytest.mark.parametrize("cfg,B", [(SMALL, 1), (SMALL, 3), (SMALL_PPG, 2), (PPG_TR, 2), (QKN_SKIP, 3)])
def test_cfm_sample_vs_oracle(cuda_dev, cfg, B):
    model, sd = build(cfg, 17, cuda_dev)
    g = torch.Generator().manual_seed(18)
    ref_len = torch.randint(40, 80, (B,), generator=g).to(cuda_dev)
    dur = torch.randint(150, 300, (B,), generator=g).to(cuda_dev)
    cond = torch.randn(B, int(ref_len.max()), cfg.mel_dim, generator=g).to(cuda_dev)
    text = torch.randint(0, cfg.text_num_embeds, (B, 30), generator=g).to(cuda_dev)
    if B > 1:
        text[1, 20:] = -1
    ppg = torch.randn(B, 100, cfg.ppg_dim, generator=g).to(cuda_dev) if cfg.use_ppg else None
    kw = dict(steps=8, cfg_strength=2.0, sway_sampling_coef=-1.0)
    # same y0 on every side: draw it once the way CFM.sample does (seeded torch RNG on the device)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    out16, traj16, aux16 = O.cfm_sample(o16, cond, text, dur, ppg=ppg, lens=ref_len, seed=3, return_aux=True, **kw)
    out32, traj32, aux32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, ppg=ppg, lens=ref_len, y0=aux16["y0"].float(),
                                        return_aux=True, **kw)
    got, traj = model.sample(cond, text, ppg, dur, lens=ref_len, seed=3, **kw)
    assert traj.shape == traj16.shape and got.dtype == torch.bfloat16
    assert torch.equal(traj[0], aux16["y0"])  # identical noise by construction
    # paste-back region is bit exact (cond copied through)
    cm = aux16["cond_mask"].expand_as(got)
    assert torch.equal(got[cm], out16[cm])
    for step in (1, kw["steps"]):
        for i in range(B):
            L = int(aux16["duration"][i])
            e_got, e_ref = rel_l2(traj[step][i, :L], traj32[step][i, :L]), rel_l2(traj16[step][i, :L], traj32[step][i, :L])
            print(f"dim={cfg.dim} B={B} step {step} sample {i}: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
            assert e_got <= K_TOL * e_ref + 5e-3, (step, i, e_got, e_ref)
    # CUDA-graph path gives the same bits as the stream path
    model.use_cuda_graph = True
    got2, _ = model.sample(cond, text, ppg, dur, lens=ref_len, seed=3, **kw)
    assert torch.equal(got2, got)


def test_full_size_f5tts_v1_base_vs_oracle(cuda_dev):
    """BASELINE configs[1] architecture at full depth and length (22 layers, dim 1024, ~940 frames), 4 utterances, 3 Euler
    steps: the oracle runs on the same GPU in fp32 and bf16, so the full-size check finishes in seconds."""
    cfg = O.DiTConfig()
    model, sd = build(cfg, 1234, cuda_dev)
    g = torch.Generator().manual_seed(0)
    B = 4
    ref_len = torch.randint(235, 421, (B,), generator=g).to(cuda_dev)
    dur = torch.randint(850, 1031, (B,), generator=g).to(cuda_dev)
    cond = torch.randn(B, int(ref_len.max()), 100, generator=g).to(cuda_dev)
    text = torch.randint(1, 2545, (B, 150), generator=g).to(cuda_dev)
    text[2, 110:] = -1
    kw = dict(steps=3, cfg_strength=2.0, sway_sampling_coef=-1.0)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    out16, traj16, aux = O.cfm_sample(o16, cond, text, dur, lens=ref_len, seed=5, return_aux=True, **kw)
    out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, lens=ref_len, y0=aux["y0"].float(), **kw)
    got, traj = model.sample(cond, text, None, dur, lens=ref_len, seed=5, **kw)
    assert torch.equal(traj[0], aux["y0"])
    for i in range(B):
        L = int(aux["duration"][i])
        e_got, e_ref = rel_l2(traj[-1][i, :L], traj32[-1][i, :L]), rel_l2(traj16[-1][i, :L], traj32[-1][i, :L])
        print(f"full size sample {i} ({L} frames): product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
        assert e_got <= K_TOL * e_ref + 5e-3, (i, e_got, e_ref)
    # size-independent property: a sample's result does not depend on what else is in the batch beyond the conv leakage
    # the reference has too: solving sample 0 alone reproduces its batched rows away from the sequence end
    got1, _ = model.sample(cond[:1, : int(ref_len[0])], text[:1], None, dur[:1], lens=ref_len[:1], seed=5, **kw)
    L0 = int(aux["duration"][0])
    assert rel_l2(got1[0, : L0 - 64].float(), got[0, : L0 - 64].float()) < 2e-2


@pytest.mark.parametrize("mode,B", [("tts", 1), ("tts", 3), ("vc", 2)])
def test_three_way_samplers_vs_oracle(cuda_dev, mode, B):
    """CFM.sample_tts / CFM.sample_vc (reference cfm.py:93-346): tripled batch, alpha_spk / alpha_txt / alpha_ppg."""
    cfg = SMALL_PPG
    model, sd = build(cfg, 27, cuda_dev)
    g = torch.Generator().manual_seed(28)
    ref_len = torch.randint(40, 80, (B,), generator=g).to(cuda_dev)
    dur = torch.randint(150, 300, (B,), generator=g).to(cuda_dev)
    cond = torch.randn(B, int(ref_len.max()), cfg.mel_dim, generator=g).to(cuda_dev)
    text = torch.randint(0, cfg.text_num_embeds, (B, 30), generator=g).to(cuda_dev)
    if B > 1:
        text[1, 20:] = -1
    ppg = torch.randn(B, 120, cfg.ppg_dim, generator=g).to(cuda_dev)
    text, ppg = (text, None) if mode == "tts" else (None, ppg)
    kw = dict(steps=6, sway_sampling_coef=-1.0)
    okw = dict(mode=mode, alpha_spk=2.5, alpha_cont=3.0, **kw)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    out16, traj16, aux = O.cfm_sample(o16, cond, text, dur, ppg=ppg, lens=ref_len, seed=3, return_aux=True, **okw)
    out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, ppg=ppg, lens=ref_len, y0=aux["y0"].float(), **okw)
    if mode == "tts":
        got, traj = model.sample_tts(cond, text, dur, lens=ref_len, seed=3, alpha_spk=2.5, alpha_txt=3.0, **kw)
    else:
        got, traj = model.sample_vc(cond, ppg, dur, lens=ref_len, seed=3, alpha_spk=2.5, alpha_ppg=3.0, **kw)
    assert traj.shape == traj16.shape and torch.equal(traj[0], aux["y0"])
    cm = aux["cond_mask"].expand_as(got)
    assert torch.equal(got[cm], out16[cm])
    for step in (1, kw["steps"]):
        for i in range(B):
            L = int(aux["duration"][i])
            e_got, e_ref = rel_l2(traj[step][i, :L], traj32[step][i, :L]), rel_l2(traj16[step][i, :L], traj32[step][i, :L])
            print(f"{mode} B={B} step {step} sample {i}: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
            assert e_got <= K_TOL * e_ref + 5e-3, (step, i, e_got, e_ref)
    model.use_cuda_graph = True
    if mode == "tts":
        got2, _ = model.sample_tts(cond, text, dur, lens=ref_len, seed=3, alpha_spk=2.5, alpha_txt=3.0, **kw)
    else:
        got2, _ = model.sample_vc(cond, ppg, dur, lens=ref_len, seed=3, alpha_spk=2.5, alpha_ppg=3.0, **kw)
    assert torch.equal(got2, got)


def test_sample_vc_needs_unmasked_text_embedder(cuda_dev):
    model, _ = build(SMALL, 7, cuda_dev)  # text_mask_padding=True: the reference raises on text=None
    with pytest.raises(ValueError):
        model.sample_vc(torch.randn(1, 30, 100, device=cuda_dev), None, 80)


def test_config3_f5e_small_ppg_batch32(cuda_dev):
    """BASELINE configs[2]: the PPG-conditioned + VQ-codebook F5E "Small" variant (configs/example.yaml: dim 768, depth 18,
    12 heads, pe_attn_head 1, text_mask_padding off, ppg 256, codebook 2 x 100), nfe 16, batch 32: `sample_vc` against the
    oracle under the bf16 bar, and the codebook indices of the model's own text / PPG embeddings bit-exact."""
    cfg = O.F5E_SMALL_PPG
    model, sd = build(cfg, 77, cuda_dev)
    g = torch.Generator().manual_seed(78)
    B = 32
    ref_len = torch.randint(90, 200, (B,), generator=g).to(cuda_dev)
    dur = torch.randint(300, 620, (B,), generator=g).to(cuda_dev)
    cond = torch.randn(B, int(ref_len.max()), cfg.mel_dim, generator=g).to(cuda_dev)
    ppg = torch.randn(B, 640, cfg.ppg_dim, generator=g).to(cuda_dev)
    kw = dict(steps=16, sway_sampling_coef=-1.0)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    okw = dict(mode="vc", alpha_spk=2.5, alpha_cont=3.0, **kw)
    out16, traj16, aux = O.cfm_sample(o16, cond, None, dur, ppg=ppg, lens=ref_len, seed=3, return_aux=True, **okw)
    out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), None, dur, ppg=ppg, lens=ref_len, y0=aux["y0"].float(), **okw)
    got, traj = model.sample_vc(cond, ppg, dur, lens=ref_len, seed=3, alpha_spk=2.5, alpha_ppg=3.0, **kw)
    assert torch.equal(traj[0], aux["y0"]) and torch.equal(aux["duration"], dur)
    worst = 0.0
    for i in range(B):
        L = int(dur[i])
        e_got, e_ref = rel_l2(traj[-1][i, :L], traj32[-1][i, :L]), rel_l2(traj16[-1][i, :L], traj32[-1][i, :L])
        worst = max(worst, e_got / (K_TOL * e_ref + 5e-3))
        assert e_got <= K_TOL * e_ref + 5e-3, (i, e_got, e_ref)
    print(f"config 3: worst product error / allowed = {worst:.2f}")
    # codebook: indices of the PPG embeddings the oracle computes in fp32, quantised by the product's kernel: bit exact
    emb = o32.ppg_embed(ppg, 640, False, B)  # [B, 640, 512] fp32
    ref = O.vq_forward(sd, emb.cpu(), groups=cfg.cb_groups, num_vars=cfg.cb_num_vars)
    xq, tg = model.transformer.quantizer.forward_idx(emb)
    assert torch.equal(tg.cpu(), ref["targets"])


def test_concurrent_callers_take_turns(cuda_dev):
    """The reference's long-form loop calls model.sample from a ThreadPoolExecutor (utils_infer.py:503-515): concurrent
    callers of ONE model must each get the bits a lone caller gets (different shapes, so a shared workspace would show)."""
    from concurrent.futures import ThreadPoolExecutor
    model, _ = build(SMALL, 17, cuda_dev)
    g = torch.Generator().manual_seed(44)
    jobs = []
    for i in range(6):
        n_ref, dur = 40 + 7 * i, 150 + 31 * i
        cond = torch.randn(1, n_ref, SMALL.mel_dim, generator=g).to(cuda_dev)
        text = torch.randint(0, SMALL.text_num_embeds, (1, 20 + i), generator=g).to(cuda_dev)
        jobs.append((cond, text, dur))

    def run(j):
        torch.cuda.set_device(cuda_dev)
        out, _ = model.sample(j[0], j[1], None, j[2], steps=4, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=5)
        torch.cuda.synchronize()
        return out

    alone = [run(j) for j in jobs]
    with ThreadPoolExecutor(max_workers=4) as ex:
        together = list(ex.map(run, jobs))
    for a, b in zip(alone, together):
        assert torch.equal(a, b)

