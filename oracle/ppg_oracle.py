"""CPU/torch restatement of the PPG extractor's front end.  TEST INFRASTRUCTURE ONLY.

Follows, by behaviour, /root/reference/src/f5_tts/ppg/: wenet/dataset/feats.py:49-83 (kaldiFbank),
wenet/transformer/cmvn.py:9-47 (GlobalCMVN), wenet/transformer/{subsampling.py:68-119, embedding.py:19-111,
attention.py:16-222, convolution.py:16-134, positionwise_feed_forward.py, encoder_layer.py:118-268, encoder.py:33-208,
403-495} (Conformer encoder, non-streaming), asr_model.py:222-244 (extract) and ppg_model.py:110-145 (lengths, mask).

Pinning: kaldi.fbank lives in torchaudio (third party, installed in this image): `kaldi_fbank` below restates its
published algorithm and tests/test_ppg_cpu.py checks it against torchaudio.compliance.kaldi.fbank itself.  The encoder
restatement is pinned against the reference's own wenet modules imported in the build container with random weights
(oracle/make_golden.py -> tests/golden/ppg_conformer.pt).  The pretrained PPG checkpoint, its train.yaml and global_cmvn
are NOT in the reference tree (ppg_model.py:11-29 loads them from outside), so the encoder configuration used here is
the code's defaults with a 12-block stack ("PPG_20ms model with 12 layers", ppg_model.py:12): parity on the real
checkpoint is unpinned.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

EPS = torch.finfo(torch.float32).eps


def povey_window(win: int) -> torch.Tensor:
    return torch.hann_window(win, periodic=False, dtype=torch.float32).pow(0.85)


def kaldi_mel_banks(num_bins=80, padded=512, sample_freq=16000.0, low_freq=20.0, high_freq=0.0) -> torch.Tensor:
    """[num_bins, padded/2] triangular banks on the kaldi mel scale (1127 ln(1 + f/700)), as kaldi.get_mel_banks without VTLN."""
    nyq = 0.5 * sample_freq
    if high_freq <= 0.0:
        high_freq += nyq
    mel = lambda f: 1127.0 * math.log(1.0 + f / 700.0)
    lo, hi = mel(low_freq), mel(high_freq)
    delta = (hi - lo) / (num_bins + 1)
    b = torch.arange(num_bins).unsqueeze(1)
    left, center, right = lo + b * delta, lo + (b + 1.0) * delta, lo + (b + 2.0) * delta
    m = (1127.0 * (1.0 + (sample_freq / padded) * torch.arange(padded // 2) / 700.0).log()).unsqueeze(0)
    return torch.max(torch.zeros(1), torch.min((m - left) / (center - left), (right - m) / (right - center)))


def kaldi_fbank(wave: torch.Tensor, sample_rate=16000, n_mels=80, frame_length_ms=25, frame_shift_ms=10,
                preemph=0.97) -> torch.Tensor:
    """feats.py:59-71 for ONE utterance: wave [L] in [-1, 1] -> [T, n_mels] log mel energies.
    kaldi.fbank defaults: snip_edges, remove_dc_offset, povey window, round_to_power_of_two, use_power, use_log_fbank."""
    x = wave.float() * (1 << 15)
    win, shift = int(sample_rate * frame_length_ms * 0.001), int(sample_rate * frame_shift_ms * 0.001)
    padded = 1 << (win - 1).bit_length()
    if x.numel() < win:
        return x.new_zeros(0, n_mels)
    T = 1 + (x.numel() - win) // shift
    fr = x.as_strided((T, win), (shift, 1)).clone()
    fr = fr - fr.mean(dim=1, keepdim=True)
    prev = torch.cat([fr[:, :1], fr[:, :-1]], dim=1)  # replicate padding on the left
    fr = fr - preemph * prev
    fr = fr * povey_window(win).to(fr.device)
    fr = F.pad(fr, (0, padded - win))
    power = torch.fft.rfft(fr).abs().pow(2.0)
    banks = F.pad(kaldi_mel_banks(n_mels, padded, float(sample_rate)), (0, 1)).to(fr.device)
    return torch.max(power @ banks.T, torch.tensor(EPS, device=fr.device)).log()


def global_cmvn(x: torch.Tensor, mean: torch.Tensor, istd: torch.Tensor) -> torch.Tensor:
    """cmvn.py:34-47."""
    return (x - mean) * istd


# ---------------------------------------------------------------------------------------------------------------
# Conformer encoder (non-streaming, eval) + content head: ASRModel.extract (asr_model.py:222-244)
# ---------------------------------------------------------------------------------------------------------------
from dataclasses import dataclass  # noqa: E402


@dataclass
class ConformerConfig:
    """encoder_conf of the PPG model.  train.yaml is not in the tree: these are the constructor defaults of
    ConformerEncoder (encoder.py:405-433) with input_layer "conv2d" (= Conv2dSubsampling2, the 20 ms PPG frame of
    ppg_model.py:12, 66) and 12 blocks."""
    input_dim: int = 80
    dim: int = 256
    heads: int = 4
    ffn: int = 2048
    blocks: int = 12
    kernel: int = 15
    cmvn: bool = True
    causal: bool = False            # convolution.py:39-45: left-only padding of kernel - 1 frames
    cnn_norm: str = "batch_norm"    # convolution.py:58-64: "batch_norm" | "layer_norm" (the streaming wenet recipes use the latter)


def make_conformer_state_dict(cfg: ConformerConfig, seed: int = 99) -> dict:
    """Random weights under the reference's own state_dict keys (ASRModel: `encoder.*`, `linear.*`)."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s, std=1.0: torch.randn(*s, generator=g) * std
    d, f, sd = cfg.dim, cfg.ffn, {}
    fdim = (cfg.input_dim - 1) // 2

    def lin(name, o, i, bias=True):
        sd[name + ".weight"] = r(o, i, std=i ** -0.5)
        if bias:
            sd[name + ".bias"] = r(o, std=0.1)

    def ln(name):
        sd[name + ".weight"] = 1.0 + r(d, std=0.1)
        sd[name + ".bias"] = r(d, std=0.1)

    if cfg.cmvn:
        sd["encoder.global_cmvn.mean"] = 10.0 + r(cfg.input_dim)
        sd["encoder.global_cmvn.istd"] = 0.3 + 0.1 * torch.rand(cfg.input_dim, generator=g)
    sd["encoder.embed.conv.0.weight"] = r(d, 1, 3, 3, std=1.0 / 3.0)
    sd["encoder.embed.conv.0.bias"] = r(d, std=0.1)
    lin("encoder.embed.out.0", d, d * fdim)
    for i in range(cfg.blocks):
        p = f"encoder.encoders.{i}."
        for ff in ("feed_forward_macaron", "feed_forward"):
            lin(p + ff + ".w_1", f, d)
            lin(p + ff + ".w_2", d, f)
        for nm in ("linear_q", "linear_k", "linear_v", "linear_out"):
            lin(p + "self_attn." + nm, d, d)
        lin(p + "self_attn.linear_pos", d, d, bias=False)
        sd[p + "self_attn.pos_bias_u"] = r(cfg.heads, d // cfg.heads, std=0.3)
        sd[p + "self_attn.pos_bias_v"] = r(cfg.heads, d // cfg.heads, std=0.3)
        sd[p + "conv_module.pointwise_conv1.weight"] = r(2 * d, d, 1, std=d ** -0.5)
        sd[p + "conv_module.pointwise_conv1.bias"] = r(2 * d, std=0.1)
        sd[p + "conv_module.depthwise_conv.weight"] = r(d, 1, cfg.kernel, std=cfg.kernel ** -0.5)
        sd[p + "conv_module.depthwise_conv.bias"] = r(d, std=0.1)
        sd[p + "conv_module.norm.weight"] = 1.0 + r(d, std=0.1)
        sd[p + "conv_module.norm.bias"] = r(d, std=0.1)
        if cfg.cnn_norm == "batch_norm":
            sd[p + "conv_module.norm.running_mean"] = r(d, std=0.1)
            sd[p + "conv_module.norm.running_var"] = 0.5 + torch.rand(d, generator=g)
        sd[p + "conv_module.pointwise_conv2.weight"] = r(d, d, 1, std=d ** -0.5)
        sd[p + "conv_module.pointwise_conv2.bias"] = r(d, std=0.1)
        for nm in ("norm_ff", "norm_mha", "norm_ff_macaron", "norm_conv", "norm_final"):
            ln(p + nm)
    ln("encoder.after_norm")
    lin("linear", d, d)
    return sd


def rel_pos_table(n: int, d: int) -> torch.Tensor:
    """embedding.py:37-43: pe[pos, 2i] = sin(pos * w_i), pe[pos, 2i+1] = cos(pos * w_i) in fp32."""
    pos = torch.arange(0, n, dtype=torch.float32).unsqueeze(1)
    div = torch.exp(torch.arange(0, d, 2, dtype=torch.float32) * -(math.log(10000.0) / d))
    pe = torch.zeros(n, d)
    pe[:, 0::2], pe[:, 1::2] = torch.sin(pos * div), torch.cos(pos * div)
    return pe


def conformer_extract(cfg: ConformerConfig, sd: dict, feats: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """ONE utterance: feats [T, input_dim] fp32 (kaldi fbank, before CMVN) -> content embedding [T', dim], T' = (T - 1) // 2.
    Every Linear / conv / LayerNorm runs in `dtype` (fp32: the reference; bf16: the precision class of the product)."""
    W = lambda k: sd[k].to(device=feats.device, dtype=dtype)
    d, H = cfg.dim, cfg.heads
    dk = d // H
    x = feats.float()
    if cfg.cmvn:
        x = (x - sd["encoder.global_cmvn.mean"].to(x.device)) * sd["encoder.global_cmvn.istd"].to(x.device)  # cmvn.py:34-47
    # Conv2dSubsampling2 (subsampling.py:81-119): Conv2d(1, d, 3, stride 2) + ReLU over (time, freq), Linear over (channel, freq)
    x = x.to(dtype)[None, None]
    x = F.relu(F.conv2d(x, W("encoder.embed.conv.0.weight"), W("encoder.embed.conv.0.bias"), stride=2))
    _, c, t, f = x.shape
    x = F.linear(x.transpose(1, 2).reshape(1, t, c * f), W("encoder.embed.out.0.weight"), W("encoder.embed.out.0.bias"))[0]
    x = x * math.sqrt(d)  # RelPositionalEncoding.forward (embedding.py:107-111); dropout off
    pos = rel_pos_table(t, d).to(device=x.device, dtype=dtype)
    ln = lambda v, name: F.layer_norm(v, (d,), W(name + ".weight"), W(name + ".bias"), 1e-5)
    swish = lambda v: v * torch.sigmoid(v)
    for i in range(cfg.blocks):
        p = f"encoder.encoders.{i}."
        # encoder_layer.py:206-213 macaron feed-forward, ff_scale 0.5
        h = ln(x, p + "norm_ff_macaron")
        h = F.linear(swish(F.linear(h, W(p + "feed_forward_macaron.w_1.weight"), W(p + "feed_forward_macaron.w_1.bias"))),
                     W(p + "feed_forward_macaron.w_2.weight"), W(p + "feed_forward_macaron.w_2.bias"))
        x = x + 0.5 * h
        # :216-236 rel-pos self attention (attention.py:179-222: no rel_shift; scores = ((q+u) k^T + (q+v) p^T) / sqrt(dk))
        h = ln(x, p + "norm_mha")
        q = F.linear(h, W(p + "self_attn.linear_q.weight"), W(p + "self_attn.linear_q.bias")).view(t, H, dk)
        k = F.linear(h, W(p + "self_attn.linear_k.weight"), W(p + "self_attn.linear_k.bias")).view(t, H, dk)
        v = F.linear(h, W(p + "self_attn.linear_v.weight"), W(p + "self_attn.linear_v.bias")).view(t, H, dk)
        pp = F.linear(pos, W(p + "self_attn.linear_pos.weight")).view(t, H, dk)
        qu = (q + W(p + "self_attn.pos_bias_u")).transpose(0, 1)
        qv = (q + W(p + "self_attn.pos_bias_v")).transpose(0, 1)
        scores = (qu @ k.permute(1, 2, 0) + qv @ pp.permute(1, 2, 0)) / math.sqrt(dk)
        a = torch.softmax(scores.float(), dim=-1).to(dtype) @ v.transpose(0, 1)
        a = a.transpose(0, 1).reshape(t, d)
        x = x + F.linear(a, W(p + "self_attn.linear_out.weight"), W(p + "self_attn.linear_out.bias"))
        # :243-248 convolution module (convolution.py:99-134): pointwise -> GLU -> depthwise k -> BatchNorm(eval) -> swish -> pointwise
        h = ln(x, p + "norm_conv").t()[None]
        if cfg.causal:  # convolution.py:104-106: the left context is padded BEFORE the biased pointwise conv, so it is glu(bias), not 0
            h = F.pad(h, (cfg.kernel - 1, 0))
        h = F.glu(F.conv1d(h, W(p + "conv_module.pointwise_conv1.weight"), W(p + "conv_module.pointwise_conv1.bias")), dim=1)
        h = F.conv1d(h, W(p + "conv_module.depthwise_conv.weight"), W(p + "conv_module.depthwise_conv.bias"),
                     padding=0 if cfg.causal else (cfg.kernel - 1) // 2, groups=d)
        if cfg.cnn_norm == "layer_norm":
            h = F.layer_norm(h.transpose(1, 2), (d,), W(p + "conv_module.norm.weight"), W(p + "conv_module.norm.bias"), 1e-5).transpose(1, 2)
        else:
            h = F.batch_norm(h, W(p + "conv_module.norm.running_mean"), W(p + "conv_module.norm.running_var"),
                             W(p + "conv_module.norm.weight"), W(p + "conv_module.norm.bias"), False, 0.0, 1e-5)
        h = F.conv1d(swish(h), W(p + "conv_module.pointwise_conv2.weight"), W(p + "conv_module.pointwise_conv2.bias"))
        x = x + h[0].t()
        # :254-263 feed-forward, final norm
        h = ln(x, p + "norm_ff")
        h = F.linear(swish(F.linear(h, W(p + "feed_forward.w_1.weight"), W(p + "feed_forward.w_1.bias"))),
                     W(p + "feed_forward.w_2.weight"), W(p + "feed_forward.w_2.bias"))
        x = ln(x + 0.5 * h, p + "norm_final")
    x = ln(x, "encoder.after_norm")  # encoder.py:203-204
    return F.linear(x, W("linear.weight"), W("linear.bias")).float()  # asr_model.py:241


def audio_to_ppg(cfg: ConformerConfig, sd: dict, wave16k: torch.Tensor, dtype=torch.float32):
    """ppg_model.py:147-168 (output_type "ppg") for one 16 kHz utterance: -> (ppg [T', dim], true_len)."""
    feats = kaldi_fbank(wave16k)
    ppg = conformer_extract(cfg, sd, feats, dtype)
    true_len = min(int(feats.shape[0] / 2), ppg.shape[0])  # :139-141 (ppg_frame_length 20 / mel_f_shift 10)
    ppg[true_len:] = 0.0
    return ppg, true_len


def get_vc_inference_prompt(metainfo, cfg: ConformerConfig, sd: dict, speed=1.0, target_sample_rate=24000, hop_length=256,
                            target_rms=0.1, infer_batch_size=1, num_buckets=200, min_secs=3, max_secs=40, dtype=torch.float32):
    """eval/utils_eval.py:245-397 with waveforms in place of file names, one utterance at a time like the reference:
    rows (utt, prompt_text, (prompt_wave, sr), gt_text, (gt_wave, gt_sr)).  The PPG comes from the 16 kHz source concatenated
    with itself (:301-320: the "prompt" half is loaded from gt_wav as well), multi-channel sources averaged after resampling."""
    import random
    import f5e_oracle as O
    prompts_all = []
    min_tokens = min_secs * target_sample_rate // hop_length
    max_tokens = max_secs * target_sample_rate // hop_length
    acc = [0] * num_buckets
    cols = [[[] for _ in range(num_buckets)] for _ in range(7)]

    def emit(b):
        mels = cols[2][b]
        T = max(m.shape[-1] for m in mels)
        padded = torch.stack([F.pad(m, (0, T - m.shape[-1]), value=0) for m in mels]).permute(0, 2, 1)
        P = max(cols[6][b])
        pp = torch.stack([F.pad(p, (0, 0, 0, P - p.shape[0]), value=0) for p in cols[5][b]])  # padded_ppg_batch (:226-242)
        prompts_all.append((cols[0][b], cols[1][b], padded, cols[3][b], cols[4][b], pp))

    for utt, prompt_text, (ref_audio, ref_sr), gt_text, (gt_audio, gt_sr) in metainfo:
        ref_audio = ref_audio.reshape(1, -1)
        ref_audio, ref_rms = O.rms_floor(ref_audio, target_rms)
        ref_audio = O.resample(ref_audio, ref_sr, target_sample_rate)
        ref_mel = O.mel_spectrogram(ref_audio).squeeze(0)
        ref_mel_len = ref_audio.shape[-1] // hop_length
        gt = gt_audio.reshape(1, -1) if gt_audio.ndim == 1 else gt_audio
        gt16 = O.resample(gt, gt_sr, 16000)
        if gt16.shape[0] > 1:
            gt16 = gt16.mean(dim=0, keepdim=True)
        full = torch.cat([gt16, gt16], dim=1)
        ppg, ppg_len = audio_to_ppg(cfg, sd, full[0], dtype)
        total = ref_mel_len + int(ref_mel_len / len(prompt_text.encode("utf-8")) * len(gt_text.encode("utf-8")) / speed)
        assert min_tokens <= total <= max_tokens
        b = math.floor((total - min_tokens) / (max_tokens - min_tokens + 1) * num_buckets)
        for c, v in zip(cols, (utt, ref_rms, ref_mel, ref_mel_len, total, ppg, ppg_len)):
            c[b].append(v)
        acc[b] += total
        if acc[b] >= infer_batch_size:
            emit(b)
            acc[b] = 0
            for c in cols:
                c[b] = []
    for b, frames in enumerate(acc):
        if frames > 0:
            emit(b)
    random.seed(666)
    random.shuffle(prompts_all)
    return prompts_all


def ppg_to_target(ppg, true_len, output_type="ppg", para_w=None, para_b=None, phn_center=None, map_mix_ratio=1.0):
    """PPGModelWapper.ppg_to_target (ppg/ppg_model.py:108-127): mask the padded frames; output_type "map" first projects
    every frame onto the phone centres: softmax(ppg @ w^T + b) @ global_phn_center, mixed with the raw PPG by map_mix_ratio."""
    T = ppg.shape[1]
    keep = (torch.arange(T, device=ppg.device)[None, :] < true_len[:, None])[:, :, None]  # ~make_pad_mask(true_len)
    if output_type == "map":
        logit = ppg @ para_w.T + para_b
        map_ppg = logit.softmax(dim=-1) @ phn_center
        ret = map_ppg if map_mix_ratio == 1.0 else ppg * (1 - map_mix_ratio) + map_ppg * map_mix_ratio
    else:
        ret = ppg
    return ret * keep
