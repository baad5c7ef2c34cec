"""GPU: the options of CFM.sample the headline path does not exercise (reference model/cfm.py:395-416, 440-441, 460-465,
478-480), the reference's default GPU dtype (fp16 modules, infer/utils_infer.py:185-194), row packing against the padded
solve, and BASELINE configs[1] at full size with the north-star tolerance form (max-abs and rel-L2 at EVERY NFE step)."""
import json
import os

import pytest
import torch

from oracle import f5e_oracle as O
from test_oracle_golden import OPTION_NAMES, load, option_case_args
from test_parity_gpu import BASE2, K_TOL, SMALL, build, rel_l2

pytestmark = pytest.mark.gpu


def _to(dev, *ts):
    return [t.to(dev) if torch.is_tensor(t) else t for t in ts]


@pytest.mark.parametrize("cname", OPTION_NAMES)
def test_sample_options_vs_oracle(cuda_dev, cname):
    """Same inputs / option values as the reference-generated fixture (tests/golden/sample_options_b3.pt), on the GPU-sized
    SMALL model: duration / cond_mask / mask bit exact vs the oracle AND vs the fixture (they depend on the inputs only),
    floating point under the bf16 bar."""
    d = load("sample_options_b3")
    c = d["cases"][cname]
    cfg = SMALL
    model, sd = build(cfg, 31, cuda_dev)
    text, dur, lens, extra = option_case_args(d, cname)
    cond, text, dur, lens = _to(cuda_dev, d["cond"], text, dur, lens)
    if "edit_mask" in extra:
        extra["edit_mask"] = extra["edit_mask"].to(cuda_dev)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    out16, traj16, aux = O.cfm_sample(o16, cond, text, dur, lens=lens, return_aux=True, **extra)
    e32 = {k: v for k, v in extra.items() if k != "seed"}
    if cname == "duplicate_test":  # y0 is blended inside; give the fp32 run the same pure noise instead of the blend
        out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, lens=lens, seed=extra["seed"], **e32)
    else:
        out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, lens=lens, y0=aux["y0"].float(), **e32)
    got, traj = model.sample(cond, text, None, dur, lens=lens, **extra)
    la = model.last_aux
    for k in ("duration", "cond_mask", "mask"):
        assert torch.equal(la[k], aux[k]), k
        assert torch.equal(la[k].cpu(), c[k]), k  # the reference's own values
    assert traj.shape == traj16.shape and traj.shape[0] == c["n_states"]
    assert torch.equal(traj[0], traj16[0])  # noise (and the duplicate_test blend): same torch ops, same bits
    assert got.shape == out16.shape == c["out"].shape and got.dtype == out16.dtype
    assert torch.isfinite(traj.float()).all()
    for i in range(cond.shape[0]):
        L = int(aux["duration"][i])
        e_got, e_ref = rel_l2(traj[-1][i, :L], traj32[-1][i, :L]), rel_l2(traj16[-1][i, :L], traj32[-1][i, :L])
        print(f"{cname} sample {i}: product-vs-fp32 {e_got:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
        assert e_got <= K_TOL * e_ref + 5e-3, (i, e_got, e_ref)
    if cname != "vocoder":  # paste-back region bit exact
        cm = aux["cond_mask"].expand_as(got)
        assert torch.equal(got[cm], out16[cm])
    else:
        assert rel_l2(got, out32) <= K_TOL * rel_l2(out16, out32) + 5e-3


def test_fp16_module(cuda_dev):
    """load_checkpoint leaves the model in fp16 on a GPU (infer/utils_infer.py:185-194, infer_cli.py's default).  The drop-in
    then draws y0 and the time grid in fp16 exactly like the reference and returns fp16, but its kernels compute in bf16 with
    fp32 accumulation (stated in DESIGN.md 2): the bar is the bf16 bar, not the fp16 one."""
    cfg = SMALL
    model, sd = build(cfg, 41, cuda_dev)
    model = model.to(torch.float16)
    g = torch.Generator().manual_seed(42)
    B = 2
    ref_len = torch.randint(40, 80, (B,), generator=g).to(cuda_dev)
    dur = torch.randint(150, 300, (B,), generator=g).to(cuda_dev)
    cond = torch.randn(B, int(ref_len.max()), cfg.mel_dim, generator=g).to(cuda_dev)
    text = torch.randint(0, cfg.text_num_embeds, (B, 30), generator=g).to(cuda_dev)
    kw = dict(steps=8, cfg_strength=2.0, sway_sampling_coef=-1.0)
    oh = O.OracleDiT(cfg, sd, torch.float16, cuda_dev)
    ob = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    outh, trajh, aux = O.cfm_sample(oh, cond, text, dur, lens=ref_len, seed=3, return_aux=True, **kw)
    y0 = aux["y0"]
    outb, trajb = O.cfm_sample(ob, cond, text, dur, lens=ref_len, y0=y0.bfloat16(), **kw)
    out32, traj32 = O.cfm_sample(o32, cond.half().float(), text, dur, lens=ref_len, y0=y0.float(), **kw)
    with pytest.warns(UserWarning, match="compute in bf16") if not getattr(type(model.transformer), "_warned_dtype", False) else _null():
        got, traj = model.sample(cond, text, None, dur, lens=ref_len, seed=3, **kw)
    assert got.dtype == torch.float16 and traj.dtype == torch.float16
    assert torch.equal(traj[0], y0)                                  # fp16 noise, bit exact
    assert torch.equal(model.last_aux["t"], aux["t"])                # fp16 time grid (sway applied in fp16), bit exact
    cm = aux["cond_mask"].expand_as(got)
    assert torch.equal(got[cm], outh[cm])
    for i in range(B):
        L = int(dur[i])
        e_got = rel_l2(traj[-1][i, :L], traj32[-1][i, :L])
        e_b, e_h = rel_l2(trajb[-1][i, :L], traj32[-1][i, :L]), rel_l2(trajh[-1][i, :L], traj32[-1][i, :L])
        print(f"fp16 module sample {i}: product {e_got:.3e}  oracle-bf16 {e_b:.3e}  oracle-fp16 {e_h:.3e} (all vs fp32)")
        assert e_got <= K_TOL * e_b + 5e-3, (i, e_got, e_b)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def test_packed_rows_tail_deviation(cuda_dev):
    """pack_valid_rows=True freezes the padded frames of shorter samples at y0 = 0, whereas the reference's evolve and leak
    through the unmasked k=31 convs into the last <= 30 valid frames.  Default (None) keeps the padded rows
    behind the valid ones and skips only the attention sub-block on them (which the reference masks to zero there): it
    must equal the padded layout BIT FOR BIT, padded frames of the trajectory included.  In the tail the default path
    stays under the bf16 bar against the oracle, and the dropping path's deviation is reported."""
    cfg = BASE2
    model, sd = build(cfg, 51, cuda_dev)
    g = torch.Generator().manual_seed(52)
    B = 3
    ref_len = torch.tensor([60, 45, 50], device=cuda_dev)
    dur = torch.tensor([400, 263, 331], device=cuda_dev)
    cond = torch.randn(B, 60, cfg.mel_dim, generator=g).to(cuda_dev)
    text = torch.randint(0, cfg.text_num_embeds, (B, 30), generator=g).to(cuda_dev)
    kw = dict(steps=8, cfg_strength=2.0, sway_sampling_coef=-1.0)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    out16, traj16, aux = O.cfm_sample(o16, cond, text, dur, lens=ref_len, seed=3, return_aux=True, **kw)
    out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, lens=ref_len, y0=aux["y0"].float(), **kw)
    res = {}
    for name, pv in (("default", None), ("padded", False), ("packed", True)):
        model.pack_valid_rows = pv
        _, traj = model.sample(cond, text, None, dur, lens=ref_len, seed=3, **kw)
        res[name] = traj[-1].float()
        res[name + "_traj"] = traj.float()
        assert model.last_aux["packed_rows"] == (pv is not False), name
        assert model.last_aux["dropped_padded_rows"] == (pv is True), name
        assert torch.isfinite(traj.float()).all()                       # no uninitialised trajectory rows
        if pv:
            pad = ~aux["mask"]
            assert (traj[1:][:, pad] == 0).all()                        # frozen padded frames are zeros, not garbage
    model.pack_valid_rows = None
    assert torch.equal(res["default_traj"], res["padded_traj"])       # every step, every row (padded frames too)
    for i in range(B):
        L = int(dur[i])
        tail = slice(max(0, L - 32), L)
        body = slice(0, max(0, L - 64))
        for name in ("padded", "packed"):
            e_t, r_t = rel_l2(res[name][i, tail], traj32[-1][i, tail]), rel_l2(traj16[-1][i, tail], traj32[-1][i, tail])
            print(f"sample {i} ({L} frames) last 32 frames, {name}: product-vs-fp32 {e_t:.3e}  oracle-bf16-vs-fp32 {r_t:.3e}")
            if name == "padded" or L == int(dur.max()):
                assert e_t <= K_TOL * r_t + 5e-3, (name, i, e_t, r_t)
        # 8 Euler steps x 2 convs x 15 taps: the frozen padding cannot reach further back than that within one evaluation;
        # the state it perturbs is local (attention mixes it in at the bf16 noise level only)
        assert rel_l2(res["packed"][i, body], res["padded"][i, body]) < 2e-2


def test_config2_exact_every_step(cuda_dev):
    """BASELINE configs[1] exactly: F5TTS_v1_Base at full depth, batch 16 x ~10 s, 32 Euler steps, CFG 2, sway -1.
    max-abs and rel-L2 of the ODE state against the fp32 oracle at EVERY NFE step, product beside the oracle run in bf16;
    the table is written to gpurun_out/parity_config2.json (committed as profiles/r02_parity_config2.json)."""
    cfg = O.DiTConfig()
    model, sd = build(cfg, 1234, cuda_dev)
    g = torch.Generator().manual_seed(0)
    B, steps = 16, 32
    ref_len = torch.randint(235, 421, (B,), generator=g).to(cuda_dev)
    dur = torch.randint(850, 1031, (B,), generator=g).to(cuda_dev)
    cond = torch.randn(B, int(ref_len.max()), 100, generator=g).to(cuda_dev)
    tlen = torch.randint(100, 201, (B,), generator=g)
    text = torch.full((B, int(tlen.max())), -1, dtype=torch.long)
    for i in range(B):
        text[i, : tlen[i]] = torch.randint(1, 2545, (int(tlen[i]),), generator=g)
    text = text.to(cuda_dev)
    kw = dict(steps=steps, cfg_strength=2.0, sway_sampling_coef=-1.0)
    o32 = O.OracleDiT(cfg, sd, torch.float32, cuda_dev)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        out16, traj16, aux = O.cfm_sample(o16, cond, text, dur, lens=ref_len, seed=5, return_aux=True, **kw)
        out32, traj32 = O.cfm_sample(o32, cond.bfloat16().float(), text, dur, lens=ref_len, y0=aux["y0"].float(), **kw)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = tf32
    got, traj = model.sample(cond, text, None, dur, lens=ref_len, seed=5, **kw)
    assert torch.equal(traj[0], aux["y0"])
    assert torch.equal(model.last_aux["duration"], aux["duration"]) and torch.equal(model.last_aux["mask"], aux["mask"])
    valid = aux["mask"].unsqueeze(-1)
    table = []
    for s in range(1, steps + 1):
        ref = traj32[s]
        dp, do = (traj[s].float() - ref) * valid, (traj16[s].float() - ref) * valid
        n = (ref * valid).norm()
        row = dict(step=s, product_rel_l2=(dp.norm() / n).item(), product_max_abs=dp.abs().max().item(),
                   oracle_bf16_rel_l2=(do.norm() / n).item(), oracle_bf16_max_abs=do.abs().max().item())
        table.append(row)
        assert row["product_rel_l2"] <= K_TOL * row["oracle_bf16_rel_l2"] + 5e-3, row
        assert row["product_max_abs"] <= K_TOL * row["oracle_bf16_max_abs"] + 5e-2, row
    worst = max(table, key=lambda r: r["product_rel_l2"])
    print(f"configs[1] exact: worst step {worst}")
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", "parity_config2.json"), "w") as f:
        json.dump(dict(config="BASELINE configs[1]: F5TTS_v1_Base depth 22, batch 16, 850-1030 frames, nfe 32, cfg 2, sway -1; "
                              "ODE state vs the fp32 oracle over the valid frames of all 16 utterances",
                       bar=f"product <= {K_TOL} x oracle-bf16 + 5e-3 (rel-L2), + 5e-2 (max-abs)", steps=table), f, indent=1)


def test_cached_graph_reuse_and_invalidation(cuda_dev):
    """use_cuda_graph: the instantiated Euler-loop graph is cached on everything it bakes in (dit.cu) and `model.py` keeps the
    solve's buffers per shape.  A replay must see THIS call's inputs, a changed weight, and another shape - each compared bit
    for bit with the stream path."""
    from f5e_tts_b200 import _lib
    cfg = SMALL
    model, sd = build(cfg, 61, cuda_dev)
    g = torch.Generator().manual_seed(62)

    def inputs(B, n_ref, n_dur, n_txt):
        ref_len = torch.randint(n_ref - 10, n_ref, (B,), generator=g).to(cuda_dev)
        dur = torch.randint(n_dur - 40, n_dur, (B,), generator=g).to(cuda_dev)
        cond = torch.randn(B, n_ref, cfg.mel_dim, generator=g).to(cuda_dev)
        text = torch.randint(0, cfg.text_num_embeds, (B, n_txt), generator=g).to(cuda_dev)
        return cond, text, dur, ref_len

    def run(graph, inp, seed):
        model.use_cuda_graph = graph
        cond, text, dur, ref_len = inp
        out, traj = model.sample(cond, text, None, dur, lens=ref_len, steps=6, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=seed)
        return out.clone(), traj.clone()

    a, b, c = inputs(2, 60, 240, 25), inputs(2, 60, 240, 25), inputs(3, 50, 180, 20)
    b = (b[0], b[1], a[2], a[3])  # same durations as `a`: the same solve shape and packing, different contents
    ref = {k: run(False, v, s) for k, (v, s) in dict(a=(a, 3), b=(b, 4), c=(c, 5)).items()}
    _lib.load().f5e_graph_cache_clear()
    counts = {}
    for key, inp, seed in (("a", a, 3), ("a", a, 3), ("b", b, 4), ("c", c, 5), ("a", a, 3), ("b", b, 4)):
        n0 = _lib.load().f5e_launch_count()
        out, traj = run(True, inp, seed)
        counts.setdefault(key, []).append(_lib.load().f5e_launch_count() - n0)
        assert torch.equal(out, ref[key][0]) and torch.equal(traj, ref[key][1]), key
    # a replayed graph accounts for the kernels it holds (bench.py's gpu_launches): capture and replays count alike
    assert len(set(counts["a"])) == 1 and counts["a"][0] > 6 * 20 and counts["b"] == counts["a"][:2], counts
    # a weight updated in place (EMA copy, hot-swapped checkpoint): plan and graphs must follow
    with torch.no_grad():
        model.transformer.transformer_blocks[0].ff.ff["2"].bias.add_(0.5)
    out_g, _ = run(True, a, 3)
    out_s, _ = run(False, a, 3)
    assert torch.equal(out_g, out_s) and not torch.equal(out_g, ref["a"][0])
    # a parent-level load_state_dict (what the reference's load_checkpoint does) restores the old weights
    missing, unexpected = model.load_state_dict({"transformer." + k: v for k, v in sd.items()}, strict=False)
    assert not unexpected
    model.to(torch.bfloat16)
    out_g, _ = run(True, a, 3)
    assert torch.equal(out_g, ref["a"][0])
    assert _lib.load().f5e_graph_cache_clear() >= 1
