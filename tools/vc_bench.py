"""Voice conversion end to end on one B200 (BASELINE configs[2] shape): 32 source utterances of ~10 s (16 kHz) -> kaldi fbank +
12-block Conformer PPG -> CFM.sample_vc on the F5E "Small" PPG + codebook model (dim 768, depth 18, 12 heads; NFE 16, three-way
guidance) -> Vocos.  Random-init weights, synthetic audio.  Prints one JSON line: generated audio seconds per second with the
source waves already on the device, and the share of each stage (CUDA events).

    python tools/vc_bench.py [--batch 32] [--nfe 16] [--iters 3]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--nfe", type=int, default=16)
    ap.add_argument("--iters", type=int, default=3)
    a = ap.parse_args()
    import ppg_oracle as PO  # weights generator only (random state dict under the reference's keys)
    from f5e_tts_b200.model import CFM, DiT
    from f5e_tts_b200.ppg import ASRModel, PPGModelWapper
    from f5e_tts_b200.vocos import Vocos
    dev = torch.device("cuda:0")
    torch.manual_seed(1234)
    ppg_config = dict(use_ppg=True, ppg_dim=256, use_transformer=False)
    cb_config = dict(use_codebook=True, num_vars=100, temp_start=2.0, temp_stop=0.5, temp_decay=0.999, groups=2, combine_groups=False,
                     weight_proj_depth=1, weight_proj_factor=1)
    dit = DiT(dim=768, depth=18, heads=12, ff_mult=2, text_dim=512, conv_layers=4, text_mask_padding=False, pe_attn_head=1,
              text_num_embeds=2545, mel_dim=100, ppg_config=ppg_config, cb_config=cb_config)
    with torch.no_grad():  # the reference zero-initialises these
        for blk in dit.transformer_blocks:
            blk.attn_norm.linear.weight.normal_(0, 0.02)
            blk.attn_norm.linear.bias.normal_(0, 0.02)
        for m in (dit.norm_out.linear, dit.proj_out):
            m.weight.normal_(0, 0.02)
            m.bias.normal_(0, 0.02)
    model = CFM(dit, ppg_config=ppg_config, cb_config=cb_config).to(dev).to(torch.bfloat16)
    model.store_trajectory = False
    voc = Vocos(seed=4321).to(dev)
    pcfg = PO.ConformerConfig(blocks=12)
    pmodel = ASRModel(pcfg.input_dim, num_blocks=12).load_state_dict(PO.make_conformer_state_dict(pcfg, seed=7)).to(dev)
    pmodel.use_cuda_graph = True
    wrap = PPGModelWapper(model=pmodel, device=dev)
    g = torch.Generator().manual_seed(0)
    B = a.batch
    src_len = torch.randint(int(8.5 * 16000), 10 * 16000 + 1, (B,), generator=g)
    src = (0.1 * torch.randn(B, int(src_len.max()), generator=g)).to(dev)
    src_len_d = src_len.to(dev)
    ref_len = torch.randint(235, 421, (B,), generator=g).to(dev)  # 2.5 - 4.5 s prompts
    gen_frames = (src_len.float() / 16000 * 24000 / 256).long().to(dev)  # VC keeps the source duration
    dur = ref_len + gen_frames
    cond = torch.randn(B, int(ref_len.max()), 100, generator=g).to(dev)
    audio_s = float(gen_frames.sum()) * 256 / 24000
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def step(timing=False):
        if timing:
            ev[0].record()
        ppg, _ = wrap.audio_to_ppg(src, 16000, lens=src_len_d)
        if timing:
            ev[1].record()
        out, _ = model.sample_vc(cond, ppg, dur, lens=ref_len, steps=a.nfe, alpha_spk=2.5, alpha_ppg=3.0, sway_sampling_coef=-1.0, seed=0)
        if timing:
            ev[2].record()
        wav = voc.decode_varlen(out, ref_len, dur)
        if timing:
            ev[3].record()
        return wav

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    acc = [0.0, 0.0, 0.0]
    for _ in range(a.iters):
        step(True)
        torch.cuda.synchronize()
        for i in range(3):
            acc[i] += ev[i].elapsed_time(ev[i + 1]) / a.iters
    tot = sum(acc)
    # the codebook quantize call (GumbelVectorQuantizer.forward, modules.py:877-950) on PPG embeddings of this batch's shape:
    # decision fused into the logit GEMM's epilogue against the unfused path (fp32 logits round trip + vq_select)
    q = dit.quantizer.float().eval()
    emb = torch.randn(B, int(gen_frames.max()) + int(ref_len.max()), 512, device=dev)
    Mq = emb.shape[0] * emb.shape[1]
    vq_bytes = Mq * (512 * 4 + 512 * 4 + 2 * 8)  # algorithmic: x read, x_q written, two int64 indices per row
    vq_ms = {}
    for name, fused in (("fused", True), ("unfused", False)):
        q.fused_decision = fused
        for _ in range(3):
            q.forward_idx(emb)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            q.forward_idx(emb)
        e1.record()
        torch.cuda.synchronize()
        vq_ms[name] = e0.elapsed_time(e1) / 10
    q.fused_decision = True
    vq = dict(rows=Mq, groups=2, num_vars=100, ms_fused=vq_ms["fused"], ms_unfused=vq_ms["unfused"], algorithmic_bytes=vq_bytes,
              gbs_fused=vq_bytes / vq_ms["fused"] / 1e6, frac_hbm_fused=vq_bytes / vq_ms["fused"] / 1e6 / 6558.1,
              note="includes the host wrapper (allocation of outputs, split3 of the fp32 input, memsets); peak = measured 6558.1 GB/s")
    print(json.dumps(dict(vq=vq, metric="voice conversion, generated audio seconds per second (PPG extraction + sample_vc + Vocos)",
                          unit="audio-s/s", value=audio_s / (tot / 1e3), ms_per_batch=tot, batch=B, nfe=a.nfe, audio_seconds=audio_s,
                          ms=dict(ppg_extraction=acc[0], sample_vc=acc[1], vocos=acc[2]), ms_per_nfe_step=acc[1] / a.nfe,
                          model="F5E Small PPG + codebook (dim 768, depth 18, 12 heads), three-way guidance (3 x batch)",
                          dtype="bf16", data="synthetic")))


if __name__ == "__main__":
    main()
