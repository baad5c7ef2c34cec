"""Generate tests/golden/*.pt by running the UNMODIFIED reference (imported from /root/reference through
oracle/shims) on seeded inputs, and print how closely the oracle restatement reproduces it.
TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference); the fixtures travel.

    python oracle/make_golden.py
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(HERE, "shims"), "/root/reference/src", os.path.dirname(HERE)]

import torch  # noqa: E402

from oracle import f5e_oracle as O  # noqa: E402
from f5_tts.model import CFM, DiT  # noqa: E402
from f5_tts.model.modules import GumbelVectorQuantizer, MelSpec  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
os.makedirs(OUT, exist_ok=True)
torch.set_num_threads(8)

SMALL = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=64, conv_layers=2, text_num_embeds=50)
SMALL_PPG = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=64, conv_layers=2, text_num_embeds=50,
                        text_mask_padding=False, pe_attn_head=1, use_ppg=True, use_codebook=True, ppg_dim=32,
                        cb_num_vars=10, cb_groups=2)


def build_reference(cfg: O.DiTConfig, seed: int):
    ppg_config = dict(use_ppg=cfg.use_ppg, ppg_dim=cfg.ppg_dim, use_transformer=cfg.ppg_transformer,
                      transformer_config=dict(nhead=cfg.ppg_nhead, dim_feedforward=cfg.ppg_ff, dropout=0.1,
                                              num_layers=cfg.ppg_layers), combined_cond_drop_prob=0.1)
    cb_config = dict(use_codebook=cfg.use_codebook, num_vars=cfg.cb_num_vars, temp_start=2.0, temp_stop=0.5,
                     temp_decay=0.999, groups=cfg.cb_groups, combine_groups=False, weight_proj_depth=1,
                     weight_proj_factor=1)
    dit = DiT(dim=cfg.dim, depth=cfg.depth, heads=cfg.heads, ff_mult=cfg.ff_mult, mel_dim=cfg.mel_dim,
              text_num_embeds=cfg.text_num_embeds, text_dim=cfg.text_dim, text_mask_padding=cfg.text_mask_padding,
              conv_layers=cfg.conv_layers, pe_attn_head=cfg.pe_attn_head, qk_norm=cfg.qk_norm,
              long_skip_connection=cfg.long_skip_connection, ppg_config=ppg_config, cb_config=cb_config)
    sd = O.make_state_dict(cfg, seed)
    missing, unexpected = dit.load_state_dict(sd, strict=False)
    missing = [k for k in missing if "num_batches_tracked" not in k and "inv_freq" not in k]
    assert not missing and not unexpected, (missing, unexpected)
    model = CFM(transformer=dit, mel_spec_kwargs=dict(), ppg_config=ppg_config, cb_config=cb_config).eval()
    return model, sd


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()


def case_sample(name, cfg, batch, seed, with_ppg=False):
    model, sd = build_reference(cfg, seed)
    g = torch.Generator().manual_seed(seed + 1)
    ref_len = torch.randint(12, 20, (batch,), generator=g)
    dur = torch.randint(40, 64, (batch,), generator=g)
    tlen = torch.randint(5, 15, (batch,), generator=g)
    T = int(ref_len.max())
    cond = torch.randn(batch, T, cfg.mel_dim, generator=g)
    text = torch.full((batch, int(tlen.max())), -1, dtype=torch.long)
    for i in range(batch):
        text[i, : tlen[i]] = torch.randint(0, cfg.text_num_embeds, (int(tlen[i]),), generator=g)
    ppg = torch.randn(batch, 30, cfg.ppg_dim, generator=g) if with_ppg else None
    kw = dict(steps=4, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=7)
    with torch.no_grad():
        out, traj = model.sample(cond=cond, text=text, ppg=ppg, duration=dur, lens=ref_len, **kw)
    dit = O.OracleDiT(cfg, sd)
    o_out, o_traj, aux = O.cfm_sample(dit, cond, text, dur, ppg=ppg, lens=ref_len, return_aux=True, **kw)
    print(f"[{name}] oracle vs reference: out rel-L2 {rel(o_out, out):.2e}  max-abs {(o_out - out).abs().max():.2e}; "
          f"traj rel-L2 {rel(o_traj, traj):.2e}")
    torch.save(dict(cfg=cfg.__dict__, seed=seed, cond=cond, text=text, ppg=ppg, duration_in=dur, lens=ref_len, kw=kw,
                    out=out, traj_last=traj[-1], traj_1=traj[1], duration=aux["duration"], mask=aux["mask"],
                    cond_mask=aux["cond_mask"]), os.path.join(OUT, name + ".pt"))
    return model, sd


def _toy_vocoder(mel):
    """Stand-in for `vocoder=` (cfm.py:478-480 only calls it on out.permute(0, 2, 1)): deterministic, shape-changing."""
    return mel.flip(1).cumsum(-1)[:, ::7].reshape(mel.shape[0], -1)


OPTION_CASES = {  # the sampler options of cfm.py:348-482 that sample_b*.pt do not exercise
    "edit_mask": dict(edit=True),
    "no_ref_audio": dict(kw=dict(no_ref_audio=True)),
    "duplicate_test": dict(kw=dict(duplicate_test=True, t_inter=0.25), steps=8),
    "vocoder": dict(vocoder=True),
    "cfg0": dict(kw=dict(cfg_strength=0.0)),
    "max_duration": dict(kw=dict(max_duration=45)),          # below two of the requested durations
    "text_longer": dict(long_text=True),                     # a text longer than the requested duration raises it
    "int_duration_no_lens": dict(int_duration=50, no_lens=True),
}


def case_sample_options(name, cfg, batch, seed):
    """One fixture holding every option case: same model / inputs, one `sample` call of the reference per option."""
    model, sd = build_reference(cfg, seed)
    dit = O.OracleDiT(cfg, sd)
    g = torch.Generator().manual_seed(seed + 1)
    ref_len = torch.randint(12, 20, (batch,), generator=g)
    dur = torch.randint(40, 64, (batch,), generator=g)
    tlen = torch.randint(5, 15, (batch,), generator=g)
    cond = torch.randn(batch, int(ref_len.max()), cfg.mel_dim, generator=g)
    text = torch.full((batch, int(tlen.max())), -1, dtype=torch.long)
    for i in range(batch):
        text[i, : tlen[i]] = torch.randint(0, cfg.text_num_embeds, (int(tlen[i]),), generator=g)
    long_text = torch.randint(0, cfg.text_num_embeds, (batch, 70), generator=g)
    long_text[1, 52:] = -1
    edit_mask = torch.ones(batch, int(ref_len.max()), dtype=torch.bool)
    edit_mask[:, 5:9] = False
    edit_mask[0, 2] = False
    fix = dict(cfg=cfg.__dict__, seed=seed, cond=cond, text=text, long_text=long_text, edit_mask=edit_mask, duration_in=dur,
               lens=ref_len, cases={})
    for cname, spec in OPTION_CASES.items():
        kw = dict(steps=spec.get("steps", 4), cfg_strength=2.0, sway_sampling_coef=-1.0, seed=7)
        kw.update(spec.get("kw", {}))
        tx = long_text if spec.get("long_text") else text
        d_in = spec.get("int_duration", dur)
        ln = None if spec.get("no_lens") else ref_len
        extra = dict(edit_mask=edit_mask) if spec.get("edit") else {}
        voc = _toy_vocoder if spec.get("vocoder") else None
        with torch.no_grad():
            out, traj = model.sample(cond=cond, text=tx, ppg=None, duration=d_in, lens=ln, vocoder=voc, **extra, **kw)
        o_out, o_traj, aux = O.cfm_sample(dit, cond, tx, d_in, lens=ln, return_aux=True, vocoder=voc, **extra, **kw)
        print(f"[{name}/{cname}] oracle vs reference: out max-abs {(o_out - out).abs().max():.2e}; traj rel-L2 {rel(o_traj, traj):.2e} "
              f"({traj.shape[0] - 1} steps); duration {aux['duration'].tolist()}")
        assert traj.shape == o_traj.shape
        c = dict(kw=kw, out=out, n_states=traj.shape[0], duration=aux["duration"], mask=aux["mask"], cond_mask=aux["cond_mask"])
        if spec.get("vocoder"):
            c["traj_last"] = traj[-1]       # `out` is the vocoded wave there
        if "duplicate_test" in kw:
            c["traj_0"] = traj[0]           # the blended start state (cfm.py:463)
        fix["cases"][cname] = c
    torch.save(fix, os.path.join(OUT, name + ".pt"))


def case_sample3(name, cfg, batch, seed, mode):
    """CFM.sample_tts / CFM.sample_vc (three-way guidance, cfm.py:93-346) on the PPG-conditioned small model."""
    model, sd = build_reference(cfg, seed)
    g = torch.Generator().manual_seed(seed + 1)
    ref_len = torch.randint(12, 20, (batch,), generator=g)
    dur = torch.randint(40, 64, (batch,), generator=g)
    tlen = torch.randint(5, 15, (batch,), generator=g)
    cond = torch.randn(batch, int(ref_len.max()), cfg.mel_dim, generator=g)
    text = torch.full((batch, int(tlen.max())), -1, dtype=torch.long)
    for i in range(batch):
        text[i, : tlen[i]] = torch.randint(0, cfg.text_num_embeds, (int(tlen[i]),), generator=g)
    ppg = torch.randn(batch, 30, cfg.ppg_dim, generator=g)
    kw = dict(steps=4, alpha_spk=2.5, sway_sampling_coef=-1.0, seed=9)
    with torch.no_grad():
        if mode == "tts":
            out, traj = model.sample_tts(cond=cond, text=text, duration=dur, lens=ref_len, alpha_txt=3.0, **kw)
            ppg = None
        else:
            out, traj = model.sample_vc(cond=cond, ppg=ppg, duration=dur, lens=ref_len, alpha_ppg=3.0, **kw)
            text = None
    dit = O.OracleDiT(cfg, sd)
    o_out, o_traj, aux = O.cfm_sample(dit, cond, text, dur, ppg=ppg, lens=ref_len, return_aux=True, mode=mode,
                                      alpha_spk=2.5, alpha_cont=3.0, steps=4, sway_sampling_coef=-1.0, seed=9)
    print(f"[{name}] oracle vs reference: out rel-L2 {rel(o_out, out):.2e}  max-abs {(o_out - out).abs().max():.2e}; "
          f"traj rel-L2 {rel(o_traj, traj):.2e}")
    torch.save(dict(cfg=cfg.__dict__, seed=seed, mode=mode, cond=cond, text=text, ppg=ppg, duration_in=dur, lens=ref_len,
                    kw=dict(steps=4, alpha_spk=2.5, alpha_cont=3.0, sway_sampling_coef=-1.0, seed=9),
                    out=out, traj_last=traj[-1], traj_1=traj[1], duration=aux["duration"], mask=aux["mask"],
                    cond_mask=aux["cond_mask"]), os.path.join(OUT, name + ".pt"))


def case_single_eval(name, cfg, seed):
    """One DiT.sample call (cond and uncond) with a per-sample mask."""
    model, sd = build_reference(cfg, seed)
    g = torch.Generator().manual_seed(seed + 2)
    B, N = 2, 48
    x = torch.randn(B, N, cfg.mel_dim, generator=g)
    cond = torch.randn(B, N, cfg.mel_dim, generator=g)
    text = torch.randint(0, cfg.text_num_embeds, (B, 11), generator=g)
    text[1, 7:] = -1
    ppg = torch.randn(B, 25, cfg.ppg_dim, generator=g) if cfg.use_ppg else None
    mask = O.lens_to_mask(torch.tensor([48, 31]))
    t = torch.tensor(0.37)
    dit = O.OracleDiT(cfg, sd)
    res = {}
    with torch.no_grad():
        for drop in (False, True):
            r = model.transformer.sample(x=x, cond=cond, text=text, ppg=ppg, time=t, mask=mask, drop_audio_cond=drop,
                                         drop_text=drop, drop_ppg=drop)
            o = dit.sample(x, cond, text, ppg, t, drop, drop, drop, mask)
            print(f"[{name}] drop={drop}: oracle vs reference rel-L2 {rel(o, r):.2e} max-abs {(o - r).abs().max():.2e}")
            res["pred_drop" if drop else "pred"] = r
        model.transformer.clear_cache()
    torch.save(dict(cfg=cfg.__dict__, seed=seed, x=x, cond=cond, text=text, ppg=ppg, mask=mask, t=t, **res),
               os.path.join(OUT, name + ".pt"))


def case_mel(name):
    g = torch.Generator().manual_seed(11)
    wav = 0.1 * torch.randn(2, 5000, generator=g)
    ref = MelSpec()(wav)
    mine = O.mel_spectrogram(wav)
    print(f"[{name}] oracle vs reference MelSpec: max-abs {(ref - mine).abs().max():.2e}")
    torch.save(dict(wav=wav, mel=ref), os.path.join(OUT, name + ".pt"))


def case_vq(name):
    torch.manual_seed(5)
    q = GumbelVectorQuantizer(dim=64, num_vars=10, temp=(2.0, 0.5, 0.999), groups=2, combine_groups=False, vq_dim=64,
                              time_first=True).eval()
    x = torch.randn(3, 17, 64)
    with torch.no_grad():
        r = q(x, produce_targets=True)
    sd = {"quantizer.vars": q.vars.data, "quantizer.weight_proj.weight": q.weight_proj.weight.data,
          "quantizer.weight_proj.bias": q.weight_proj.bias.data}
    o = O.vq_forward(sd, x, groups=2, num_vars=10)
    print(f"[{name}] targets equal: {torch.equal(o['targets'], r['targets'])}; x max-abs {(o['x'] - r['x']).abs().max():.2e}; "
          f"ppl {float(o['code_perplexity'] - r['code_perplexity']):.1e} {float(o['prob_perplexity'] - r['prob_perplexity']):.1e}")
    torch.save(dict(sd=sd, x=x, xq=r["x"], targets=r["targets"], code_perplexity=r["code_perplexity"],
                    prob_perplexity=r["prob_perplexity"]), os.path.join(OUT, name + ".pt"))


def case_resample(name):
    """torchaudio.transforms.Resample (the reference's resampler, eval/utils_eval.py:125-127) on seeded noise."""
    import torchaudio
    g = torch.Generator().manual_seed(21)
    cases = []
    for orig, new, n in [(16000, 24000, 9001), (44100, 24000, 12345), (48000, 24000, 8000), (22050, 24000, 7001)]:
        x = 0.1 * torch.randn(2, n, generator=g)
        y = torchaudio.transforms.Resample(orig, new)(x)
        print(f"[{name}] {orig}->{new}: oracle vs torchaudio max-abs {(O.resample(x, orig, new) - y).abs().max():.2e}")
        cases.append(dict(orig=orig, new=new, x=x, y=y))
    torch.save(cases, os.path.join(OUT, name + ".pt"))


def case_chunk_text(name):
    """chunk_text of the reference (infer/utils_infer.py:70-98), executed from its source without importing the module
    (the module pulls pydub / transformers pipelines at import time)."""
    import ast
    import re
    src = open("/root/reference/src/f5_tts/infer/utils_infer.py").read()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "chunk_text")
    ns = {"re": re}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "utils_infer.py", "exec"), ns)
    texts = [
        "Hello there. This is a test of the chunker, which splits on punctuation; it keeps going: really! Does it? Yes.",
        "短句。这是一个比较长的中文句子，用来测试全角标点的切分；没有空格！对吗？是的。",
        "No punctuation at all just a very long run of words that never ends " * 5,
        "Mixed 中文 and English, with commas, and periods. 还有句号。And more!",
        "",
        "One.",
    ]
    rows = [dict(text=t, max_chars=m, chunks=ns["chunk_text"](t, max_chars=m)) for t in texts for m in (135, 40)]
    torch.save(rows, os.path.join(OUT, name + ".pt"))
    print(f"[{name}] {len(rows)} cases")


def case_ppg_conformer(name, causal=False, cnn_norm="batch_norm"):
    """The reference's own wenet ConformerEncoder (ppg/wenet/transformer/encoder.py:403-495) + the content Linear of
    ASRModel (asr_model.py:78-79, 241), built the way init_asr_model does (asr_model.py:814-836; the decoder / CTC heads
    are not on the extract path) with random weights, on the reference's kaldiFbank features of seeded noise."""
    from f5_tts.ppg.wenet.dataset.feats import kaldiFbank
    from f5_tts.ppg.wenet.transformer.cmvn import GlobalCMVN
    from f5_tts.ppg.wenet.transformer.encoder import ConformerEncoder
    from oracle import ppg_oracle as PO
    cfg = PO.ConformerConfig(blocks=3, causal=causal, cnn_norm=cnn_norm)
    sd = PO.make_conformer_state_dict(cfg, seed=99)
    enc = ConformerEncoder(cfg.input_dim, global_cmvn=GlobalCMVN(sd["encoder.global_cmvn.mean"], sd["encoder.global_cmvn.istd"]),
                           output_size=cfg.dim, attention_heads=cfg.heads, linear_units=cfg.ffn, num_blocks=cfg.blocks,
                           input_layer="conv2d", cnn_module_kernel=cfg.kernel, causal=causal, cnn_module_norm=cnn_norm).eval()
    esd = {k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")}
    missing, unexpected = enc.load_state_dict(esd, strict=False)
    missing = [k for k in missing if "num_batches_tracked" not in k and "linear_xs_embs" not in k and "concat_linear" not in k]
    assert not missing and not unexpected, (missing, unexpected)
    linear = torch.nn.Linear(cfg.dim, cfg.dim)
    linear.load_state_dict({"weight": sd["linear.weight"], "bias": sd["linear.bias"]})
    g = torch.Generator().manual_seed(31)
    wave = 0.1 * torch.randn(1, 20000, generator=g) * torch.linspace(0.2, 1.0, 20000)
    with torch.no_grad():
        feats, feats_len = kaldiFbank().eval()(wave)
        out, mask = enc(feats, feats_len, embs=None, decoding_chunk_size=-1)
        ref = linear(out)[0]
    got_feats = PO.kaldi_fbank(wave[0])
    got = PO.conformer_extract(cfg, sd, got_feats)
    print(f"{name}: fbank max abs {float((got_feats - feats[0]).abs().max()):.2e}; extract rel-L2 {rel(got, ref):.2e} "
          f"shape {tuple(ref.shape)} mask frames {int(mask.sum())}")
    torch.save(dict(blocks=cfg.blocks, seed=99, causal=causal, cnn_norm=cnn_norm, wave=wave[0], feats=feats[0], ppg=ref),
               os.path.join(OUT, name + ".pt"))


def case_ppg_map(name):
    """PPGModelWapper.ppg_to_target with output_type "map" (ppg_model.py:108-127) of the reference itself; the wrapper's
    __init__ loads a checkpoint and two tables from paths outside the tree, so the object is built without it."""
    from f5_tts.ppg.ppg_model import PPGModelWapper
    from oracle import ppg_oracle as PO

    g = torch.Generator().manual_seed(321)
    B, T, dim, V = 2, 23, 64, 77
    ppg = torch.randn(B, T, dim, generator=g)
    true_len = torch.tensor([23, 17])
    w, b = torch.randn(V, dim, generator=g) / 4, torch.randn(V, generator=g)
    center = torch.randn(V, dim, generator=g)
    out = {}
    for ratio in (1.0, 0.35):
        ref = object.__new__(PPGModelWapper)
        ref.output_type, ref.map_mix_ratio = "map", ratio
        ref.global_phn_center, ref.para_softmax = center, dict(w=w, b=b)
        out[ratio] = ref.ppg_to_target(ppg, true_len)
        mine = PO.ppg_to_target(ppg, true_len, "map", w, b, center, ratio)
        print(f"[{name}] ratio {ratio}: oracle vs reference max-abs {(mine - out[ratio]).abs().max():.2e}")
    torch.save(dict(ppg=ppg, true_len=true_len, w=w, b=b, center=center, out_1=out[1.0], out_035=out[0.35]),
               os.path.join(OUT, name + ".pt"))


SMALL_QKN = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=64, conv_layers=2, text_num_embeds=50, qk_norm="rms_norm",
                        long_skip_connection=True, pe_attn_head=2)

if __name__ == "__main__":
    case_ppg_map("ppg_map")
    case_single_eval("dit_eval_small_qknorm_skip", SMALL_QKN, 103)
    case_mel("melspec")
    case_vq("vq")
    case_single_eval("dit_eval_small", SMALL, 100)
    case_single_eval("dit_eval_small_ppg", SMALL_PPG, 101)
    case_sample("sample_b1", SMALL, 1, 200)
    case_sample("sample_b3", SMALL, 3, 201)
    case_sample("sample_b2_ppg", SMALL_PPG, 2, 202, with_ppg=True)
    case_sample_options("sample_options_b3", SMALL, 3, 205)
    case_sample3("sample_tts_b2", SMALL_PPG, 2, 203, "tts")
    case_sample3("sample_vc_b2", SMALL_PPG, 2, 204, "vc")
    SMALL_PPG_TR = O.DiTConfig(dim=256, depth=2, heads=4, text_dim=64, conv_layers=2, text_num_embeds=50,
                               text_mask_padding=False, pe_attn_head=1, use_ppg=True, ppg_dim=32, ppg_transformer=True,
                               ppg_nhead=2, ppg_ff=48, ppg_layers=2)
    case_single_eval("dit_eval_small_ppg_transformer", SMALL_PPG_TR, 102)
    case_resample("resample")
    case_chunk_text("chunk_text")
    case_ppg_conformer("ppg_conformer")
    case_ppg_conformer("ppg_conformer_causal_ln", causal=True, cnn_norm="layer_norm")  # the streaming wenet recipe's conv module
