"""One small product-only pass over every kernel family of the hot path, sized to run under
`compute-sanitizer --tool memcheck` (ONE tool per gpurun call, B200_PROFILING.md) in a few minutes:

  * CFM.sample on a 2-block F5TTS_v1_Base-width DiT, ragged batch 4 x ~1000 frames with CFG 2: the rows are enough for the
    folded-LayerNorm GEMMs, the CTA-pair tiles, the statistics-taking FF2 epilogue and the split row packing; narrow last key
    block in attention; text embedder kernels; fused CFG + Euler
  * the same solve through the cached CUDA graph, twice (capture, then replay)
  * a batch-1 solve (no mask, single-CTA tiles)
  * a qk_norm + long-skip + partial-RoPE small model
  * MelSpec, Vocos decode / decode_varlen, the fused and the unfused VQ decision

No oracle, no assertions beyond finiteness: the point is the sanitizer's verdict.  Usage:
  python tools/sanitize_case.py && compute-sanitizer --tool memcheck --error-exitcode 7 python tools/sanitize_case.py
(Round 2: compute-sanitizer is CLOSED on this GPU pool - the call is refused - so only the plain run exists: 803 launches,
every result finite.  Out-of-bounds protection therefore rests on the F5E_REQUIRE argument checks of every entry point, TMA's
hardware clipping of partial tiles, and the parity tests at ragged / odd sizes.)
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from f5e_tts_b200 import _lib  # noqa: E402
from f5e_tts_b200.model import CFM, DiT  # noqa: E402
from f5e_tts_b200.frontend import MelSpec  # noqa: E402
from f5e_tts_b200.quantizer import GumbelVectorQuantizer  # noqa: E402
from f5e_tts_b200.vocos import Vocos  # noqa: E402


def randomise(module, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for p in module.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (0.02 if p.dim() > 1 else 0.1))


def solve(model, dev, B, lo, hi, seed, **kw):
    g = torch.Generator().manual_seed(seed)
    ref_len = torch.randint(40, 80, (B,), generator=g).to(dev)
    dur = torch.randint(lo, hi, (B,), generator=g).to(dev)
    cond = torch.randn(B, int(ref_len.max()), 100, generator=g).to(dev)
    text = torch.randint(0, 200, (B, 40), generator=g).to(dev)
    out, traj = model.sample(cond, text, None, dur, lens=ref_len, seed=3, steps=2, cfg_strength=2.0, sway_sampling_coef=-1.0, **kw)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all() and torch.isfinite(traj.float()).all()
    return out


def main():
    dev = torch.device("cuda:0")
    _lib.require_device()
    n0 = _lib.load().f5e_launch_count()

    dit = DiT(dim=1024, depth=2, heads=16, ff_mult=2, mel_dim=100, text_num_embeds=256, text_dim=512, conv_layers=4)
    randomise(dit, 1)
    model = CFM(dit).to(dev).to(torch.bfloat16)
    out = solve(model, dev, 4, 900, 1031, 11)
    assert model.last_aux["packed_rows"] and not model.last_aux["dropped_padded_rows"]
    for pv in (False, True):
        model.pack_valid_rows = pv
        solve(model, dev, 4, 900, 1031, 11)
    model.pack_valid_rows = None
    model.use_cuda_graph = True
    solve(model, dev, 4, 900, 1031, 11)
    solve(model, dev, 4, 900, 1031, 11)
    solve(model, dev, 1, 300, 301, 12)
    solve(model, dev, 1, 300, 301, 12)
    model.use_cuda_graph = False
    _lib.load().f5e_graph_cache_clear()
    print("dit: ok")

    dit2 = DiT(dim=512, depth=2, heads=8, ff_mult=2, mel_dim=100, text_num_embeds=256, text_dim=256, conv_layers=2,
               qk_norm="rms_norm", long_skip_connection=True, pe_attn_head=1, text_mask_padding=False)
    randomise(dit2, 2)
    solve(CFM(dit2).to(dev).to(torch.bfloat16), dev, 3, 150, 400, 13)
    print("dit (qk_norm, long skip, pe_attn_head=1): ok")

    wav = 0.1 * torch.randn(2, 24000, generator=torch.Generator().manual_seed(5)).to(dev)
    mel = MelSpec()(wav)
    assert torch.isfinite(mel).all()
    voc = Vocos(seed=4321).to(dev)
    w = voc.decode(mel)
    assert torch.isfinite(w).all()
    start = torch.tensor([60, 45, 70, 50], device=dev)
    end = torch.tensor([400, 263, 331, 380], device=dev)
    w2 = voc.decode_varlen(out.float()[:, :400].contiguous(), start, end)
    assert torch.isfinite(w2).all()
    torch.cuda.synchronize()
    print("melspec / vocos: ok")

    torch.manual_seed(21)
    q = GumbelVectorQuantizer(dim=512, num_vars=100, temp=(2.0, 0.5, 0.999), groups=2, combine_groups=False, vq_dim=512,
                              time_first=True).eval().to(dev)
    x = torch.randn(4, 333, 512, generator=torch.Generator().manual_seed(6)).to(dev)
    r = q(x, produce_targets=True)
    q.fused_decision = False
    r0 = q(x, produce_targets=True)
    assert torch.equal(r["targets"], r0["targets"])
    torch.cuda.synchronize()
    print("vq: ok")
    print("launches:", _lib.load().f5e_launch_count() - n0)


if __name__ == "__main__":
    main()
