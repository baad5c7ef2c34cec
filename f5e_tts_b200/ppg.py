"""Front end of the PPG extractor on the GPU (first half): the reference's `kaldiFbank`
(ppg/wenet/dataset/feats.py:49-83) and `GlobalCMVN` (ppg/wenet/transformer/cmvn.py:9-47) as ONE kernel
(`f5e_kaldi_fbank`), plus the length / padding-mask logic of `PPGModelWapper` (ppg/ppg_model.py:110-168).

Same names and argument meaning as the reference.  `kaldiFbank.forward` additionally takes `lens` (a padded batch of
utterances of different length: every utterance gets the frames a batch-1 call would give it, the rest are zeros) and
`cmvn` (fuses the encoder's GlobalCMVN into the same pass).  There is no CPU path.
"""
from __future__ import annotations

import math
import os

import torch
from torch import nn

from . import _lib
from ._lib import check, ptr, stream_ptr

_TABLES: dict = {}


def povey_window(win_length: int) -> torch.Tensor:
    """torchaudio.compliance.kaldi window_type='povey': hann(periodic=False) ** 0.85, fp32."""
    return torch.hann_window(win_length, periodic=False, dtype=torch.float32).pow(0.85)


def kaldi_mel_banks(num_bins=80, padded=512, sample_freq=16000.0, low_freq=20.0, high_freq=0.0) -> torch.Tensor:
    """fp32 [num_bins, padded // 2 + 1]: the algorithm of torchaudio.compliance.kaldi.get_mel_banks (no VTLN) with the
    zero Nyquist column kaldi.fbank appends; same torch fp32 operations in the same order, so the table matches bit for bit."""
    num_fft_bins = padded // 2
    nyquist = 0.5 * sample_freq
    if high_freq <= 0.0:
        high_freq += nyquist
    fft_bin_width = sample_freq / padded
    mel = lambda f: 1127.0 * math.log(1.0 + f / 700.0)
    mel_low, mel_high = mel(low_freq), mel(high_freq)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = torch.arange(num_bins).unsqueeze(1)
    left, center, right = mel_low + b * delta, mel_low + (b + 1.0) * delta, mel_low + (b + 2.0) * delta
    m = (1127.0 * (1.0 + fft_bin_width * torch.arange(num_fft_bins) / 700.0).log()).unsqueeze(0)
    up, down = (m - left) / (center - left), (right - m) / (right - center)
    bins = torch.max(torch.zeros(1), torch.min(up, down))
    return torch.nn.functional.pad(bins, (0, 1), mode="constant", value=0.0).float().contiguous()


def _fbank_tables(device, n_mels, win, n_fft, sample_rate):
    key = (str(device), n_mels, win, n_fft, sample_rate)
    if key not in _TABLES:
        fb = kaldi_mel_banks(n_mels, n_fft, float(sample_rate))
        nz = fb > 0
        n_bin = fb.shape[1]
        lo = torch.where(nz.any(1), nz.float().argmax(1), torch.zeros(n_mels, dtype=torch.long))
        hi = torch.where(nz.any(1), n_bin - nz.flip(1).float().argmax(1), torch.zeros(n_mels, dtype=torch.long))
        ang = -2.0 * math.pi * torch.arange(n_fft // 2, dtype=torch.float64) / n_fft
        tw = torch.stack((ang.cos(), ang.sin()), dim=-1).to(torch.float32)
        _TABLES[key] = tuple(t.to(device).contiguous() for t in
                             (povey_window(win), tw, fb, lo.to(torch.int32), hi.to(torch.int32)))
    return _TABLES[key]


class GlobalCMVN(nn.Module):
    """ppg/wenet/transformer/cmvn.py:9-47: (x - mean) * istd.  Used fused (`kaldiFbank.forward(cmvn=...)`)."""

    def __init__(self, mean: torch.Tensor, istd: torch.Tensor, norm_var: bool = True):
        super().__init__()
        assert mean.shape == istd.shape
        self.norm_var = norm_var
        self.register_buffer("mean", mean.float())
        self.register_buffer("istd", istd.float())


class kaldiFbank(nn.Module):
    """ppg/wenet/dataset/feats.py:49-83.  forward(x [B, L] in [-1, 1]) -> (feats fp32 [B, T, n_mels], speech_lengths)."""

    def __init__(self, sample_rate=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80, spec_mask_time=(5, 10),
                 spec_mask_freq=(5, 10)):
        super().__init__()
        self.f_length = int(win_length / sample_rate * 1000)  # ms, as the reference passes them on
        self.f_shift = int(hop_length / sample_rate * 1000)
        self.sample_rate, self.n_fft, self.n_mels = sample_rate, n_fft, n_mels
        if n_fft != 512:
            raise NotImplementedError("f5e_tts_b200.kaldiFbank: the kernel is specialised for the 512-point FFT")

    @torch.no_grad()
    def forward(self, x: torch.Tensor, is_spec_aug=(), *, lens: torch.Tensor | None = None, cmvn: GlobalCMVN | None = None):
        if len(is_spec_aug) != 0:
            raise NotImplementedError("f5e_tts_b200.kaldiFbank: SpecAugment is a training-time option (inference path only)")
        if not x.is_cuda:
            raise RuntimeError("f5e_tts_b200.kaldiFbank has no CPU path: pass a CUDA tensor")
        assert x.dim() == 2
        _lib.require_device()
        x = x.float()
        if x.stride(1) != 1:
            x = x.contiguous()
        B, L = x.shape
        # kaldi: window_size = int(sample_frequency * frame_length * 0.001), padded to the next power of two
        win, shift = int(self.sample_rate * self.f_length * 0.001), int(self.sample_rate * self.f_shift * 0.001)
        if L < win:
            raise ValueError(f"kaldiFbank: {L} samples are shorter than one {win}-sample frame")
        t_max = 1 + (L - win) // shift
        window, tw, fb, lo, hi = _fbank_tables(x.device, self.n_mels, win, self.n_fft, self.sample_rate)
        lens32 = None if lens is None else lens.to(device=x.device, dtype=torch.int32).contiguous()
        out = torch.empty(B, t_max, self.n_mels, device=x.device, dtype=torch.float32)
        mean = istd = None
        if cmvn is not None:
            mean = cmvn.mean.to(x.device).contiguous()
            istd = (cmvn.istd if cmvn.norm_var else torch.ones_like(cmvn.istd)).to(x.device).contiguous()
        f32, i64 = _lib.C.c_float, _lib.C.c_int64
        check(_lib.load().f5e_kaldi_fbank(ptr(x), i64(x.stride(0)), ptr(lens32), B, L, f32(float(1 << 15)), f32(0.97), win, shift,
                                          ptr(window), ptr(tw), ptr(fb), ptr(lo), ptr(hi), self.n_mels, ptr(mean), ptr(istd),
                                          ptr(out), i64(out.stride(0)), t_max, stream_ptr()), "f5e_kaldi_fbank")
        if lens is None:
            speech_lengths = torch.tensor([t_max])  # feats.py:82: one length for the whole (equal-length) batch
        else:
            n = lens.to(torch.long)
            speech_lengths = torch.where(n >= win, 1 + (n - win) // shift, torch.zeros_like(n))
        return out, speech_lengths


def make_pad_mask(lengths: torch.Tensor, max_len: int = 0) -> torch.Tensor:
    """ppg/ppg_model.py:31-57: True on the padded part."""
    max_len = max_len if max_len > 0 else int(lengths.max())
    return torch.arange(max_len, device=lengths.device).unsqueeze(0) >= lengths.unsqueeze(-1)


def ppg_true_len(mel_lens: torch.Tensor, ppg_len: int, ppg_frame_length=20, mel_f_shift=10) -> torch.Tensor:
    """ppg/ppg_model.py:139-141: frames of the (sub-sampled) PPG that belong to each utterance."""
    return (mel_lens / (ppg_frame_length / mel_f_shift)).long().clamp(max=ppg_len)


# ---------------------------------------------------------------------------------------------------------------
# Conformer encoder + content head (ASRModel.extract, ppg/asr_model.py:222-244) and the wrapper (ppg/ppg_model.py:59-168)
# ---------------------------------------------------------------------------------------------------------------
from . import ops  # noqa: E402

_ENCODER_DEFAULTS = dict(output_size=256, attention_heads=4, linear_units=2048, num_blocks=6, input_layer="conv2d",
                         pos_enc_layer_type="rel_pos", normalize_before=True, concat_after=False, macaron_style=True,
                         activation_type="swish", use_cnn_module=True, cnn_module_kernel=15, causal=False,
                         cnn_module_norm="batch_norm", use_emb=False)


class ASRModel:
    """The part of the reference's ASRModel that `extract` runs (asr_model.py:222-244, non-streaming): GlobalCMVN,
    Conv2dSubsampling2, RelPositionalEncoding, `num_blocks` ConformerEncoderLayers, after_norm and the content Linear.
    `load_state_dict` takes the reference checkpoint's keys (`encoder.*`, `linear.*`; decoder / CTC / CE heads are not on
    this path and are ignored).  Activations bf16, statistics and accumulation fp32; no CPU path."""

    def __init__(self, input_dim: int = 80, **encoder_conf):
        conf = dict(_ENCODER_DEFAULTS)
        unknown = set(encoder_conf) - set(conf) - {"dropout_rate", "positional_dropout_rate", "attention_dropout_rate",
                                                    "static_chunk_size", "use_dynamic_chunk", "use_dynamic_left_chunk",
                                                    "selfattention_layer_type", "positionwise_conv_kernel_size", "emb_dim"}
        if unknown:
            raise TypeError(f"ASRModel: unknown encoder_conf keys {sorted(unknown)}")
        conf.update({k: v for k, v in encoder_conf.items() if k in conf})
        for key, want in (("input_layer", "conv2d"), ("pos_enc_layer_type", "rel_pos"), ("normalize_before", True),
                          ("concat_after", False), ("macaron_style", True), ("activation_type", "swish"),
                          ("use_cnn_module", True), ("use_emb", False)):
            if conf[key] != want:
                raise NotImplementedError(f"f5e_tts_b200.ASRModel: {key}={conf[key]!r} is not implemented (only {want!r})")
        if conf["cnn_module_norm"] not in ("batch_norm", "layer_norm"):
            raise ValueError(f"cnn_module_norm must be 'batch_norm' or 'layer_norm', not {conf['cnn_module_norm']!r}")
        self.conf, self.input_dim = conf, input_dim
        self.dim, self.heads = conf["output_size"], conf["attention_heads"]
        if self.dim % self.heads or self.dim // self.heads != 64 or self.dim % 256:
            raise NotImplementedError("f5e_tts_b200.ASRModel: the attention kernel needs output_size / attention_heads == 64 "
                                      "and the norm kernel output_size % 256 == 0")
        self.sd: dict | None = None
        self.device = None
        self._plan = None
        self._pos: dict = {}
        self._graphs: dict = {}
        self.use_cuda_graph = False  # extension: replay one captured graph per (batch, frames) shape
        self.max_cached_graphs = 8

    def load_state_dict(self, state_dict: dict, strict: bool = False):
        need = ["encoder.embed.conv.0.weight", "encoder.embed.out.0.weight", "encoder.after_norm.weight", "linear.weight"]
        need += [f"encoder.encoders.{self.conf['num_blocks'] - 1}.norm_final.weight"]
        missing = [k for k in need if k not in state_dict]
        if missing:
            raise KeyError(f"ASRModel.load_state_dict: missing {missing}")
        self.sd = {k: v.detach().float() for k, v in state_dict.items() if k.startswith("encoder.") or k.startswith("linear.")}
        self._plan, self._pos, self._graphs = None, {}, {}
        return self

    def to(self, device):
        self.device = torch.device(device)
        self._plan, self._pos, self._graphs = None, {}, {}
        return self

    def eval(self):
        return self

    @property
    def global_cmvn(self) -> GlobalCMVN | None:
        if self.sd is None or "encoder.global_cmvn.mean" not in self.sd:
            return None
        return GlobalCMVN(self.sd["encoder.global_cmvn.mean"], self.sd["encoder.global_cmvn.istd"])

    # ---- plan: device copies of the weights in the layouts the kernels take -------------------------------------
    def _get_plan(self):
        if self._plan is not None:
            return self._plan
        if self.sd is None or self.device is None or self.device.type != "cuda":
            raise RuntimeError("f5e_tts_b200.ASRModel: load_state_dict() and .to(<cuda device>) first; there is no CPU path")
        _lib.require_device()
        sd, dev, d = self.sd, self.device, self.dim
        f32 = lambda t: t.to(dev, torch.float32).contiguous()
        b16 = lambda t: t.to(dev, torch.bfloat16).contiguous()
        fo = (self.input_dim - 3) // 2 + 1
        xs = math.sqrt(d)  # RelPositionalEncoding.xscale, folded into the sub-sampling Linear
        w_out = sd["encoder.embed.out.0.weight"].view(d, d, fo).permute(0, 2, 1).reshape(d, fo * d)  # (c, f) -> (f, c) columns
        P = dict(fo=fo, conv_w=f32(sd["encoder.embed.conv.0.weight"].view(d, 9)), conv_b=f32(sd["encoder.embed.conv.0.bias"]),
                 w_out=b16(w_out * xs), b_out=f32(sd["encoder.embed.out.0.bias"] * xs),
                 an_w=f32(sd["encoder.after_norm.weight"]), an_b=f32(sd["encoder.after_norm.bias"]),
                 w_lin=b16(sd["linear.weight"]), b_lin=f32(sd["linear.bias"]), layers=[])
        taps = self.conf["cnn_module_kernel"]
        for i in range(self.conf["num_blocks"]):
            p = f"encoder.encoders.{i}."
            g = lambda k: sd[p + k]
            ly = {}
            for nm in ("norm_ff_macaron", "norm_mha", "norm_conv", "norm_ff", "norm_final"):
                ly[nm] = (f32(g(nm + ".weight")), f32(g(nm + ".bias")))
            for nm, tag in (("feed_forward_macaron", "m"), ("feed_forward", "f")):
                ly[tag + "w1"], ly[tag + "b1"] = b16(g(nm + ".w_1.weight")), f32(g(nm + ".w_1.bias"))
                ly[tag + "w2"], ly[tag + "b2"] = b16(0.5 * g(nm + ".w_2.weight")), f32(0.5 * g(nm + ".w_2.bias"))  # ff_scale
            ly["wqkv"] = b16(torch.cat([g("self_attn.linear_q.weight"), g("self_attn.linear_k.weight"), g("self_attn.linear_v.weight")]))
            ly["bqkv"] = f32(torch.cat([g("self_attn.linear_q.bias"), g("self_attn.linear_k.bias"), g("self_attn.linear_v.bias")]))
            ly["wo"], ly["bo"] = b16(g("self_attn.linear_out.weight")), f32(g("self_attn.linear_out.bias"))
            ly["wpos"] = g("self_attn.linear_pos.weight")
            ly["u"], ly["v"] = f32(g("self_attn.pos_bias_u")), f32(g("self_attn.pos_bias_v"))
            ly["wpw1"], ly["bpw1"] = b16(g("conv_module.pointwise_conv1.weight").view(2 * d, d)), f32(g("conv_module.pointwise_conv1.bias"))
            wd, bd = g("conv_module.depthwise_conv.weight").view(d, taps), g("conv_module.depthwise_conv.bias")
            if self.conf["cnn_module_norm"] == "batch_norm":  # eval BatchNorm1d is affine per channel: folded into the conv
                s = g("conv_module.norm.weight") / torch.sqrt(g("conv_module.norm.running_var") + 1e-5)
                wd, bd = wd * s[:, None], (bd - g("conv_module.norm.running_mean")) * s + g("conv_module.norm.bias")
                ly["ln_w"] = ly["ln_b"] = None
            else:  # LayerNorm over channels: applied inside glu_dwconv
                ly["ln_w"], ly["ln_b"] = f32(g("conv_module.norm.weight")), f32(g("conv_module.norm.bias"))
            ly["wdw"], ly["bdw"] = f32(wd), f32(bd)
            ly["pad"] = None
            if self.conf["causal"]:  # convolution.py:104-113: the left context passes the biased pointwise conv and the GLU
                bq = g("conv_module.pointwise_conv1.bias").to(torch.bfloat16).float()  # what the GEMM writes for a zero input row
                ly["pad"] = f32(bq[:d] * torch.sigmoid(bq[d:]))
            ly["wpw2"], ly["bpw2"] = b16(g("conv_module.pointwise_conv2.weight").view(d, d)), f32(g("conv_module.pointwise_conv2.bias"))
            P["layers"].append(ly)
        self._plan = P
        return P

    def _pos_tables(self, n_pad: int):
        """linear_pos(pos_emb[:n_pad]) of every layer (attention.py:199-200): input independent, fp32 [n_pad, dim]."""
        if n_pad not in self._pos:
            pos = torch.arange(0, n_pad, dtype=torch.float32).unsqueeze(1)
            div = torch.exp(torch.arange(0, self.dim, 2, dtype=torch.float32) * -(math.log(10000.0) / self.dim))
            pe = torch.zeros(n_pad, self.dim)
            pe[:, 0::2], pe[:, 1::2] = torch.sin(pos * div), torch.cos(pos * div)  # embedding.py:37-43
            pe = pe.to(self.device)
            self._pos[n_pad] = [(pe @ ly["wpos"].to(self.device).t()).contiguous() for ly in self._get_plan()["layers"]]
        return self._pos[n_pad]

    @torch.no_grad()
    def extract(self, speech: torch.Tensor, speech_lengths: torch.Tensor, stream: bool = False, *, cmvn_applied: bool = False):
        """asr_model.py:222-244.  speech fp32 [B, T, input_dim] fbank features, speech_lengths [B] (or [1] for all) frames.
        -> (content embedding fp32 [B, T', dim], None); every utterance is processed as a batch-1 call would process it,
        rows beyond its own (len - 1) // 2 frames are zeros.  The CE logits of the reference are not on the PPG path."""
        if stream:
            raise NotImplementedError("f5e_tts_b200.ASRModel.extract: chunk-by-chunk streaming is not implemented")
        P = self._get_plan()
        lib, dev, d, H = _lib.load(), self.device, self.dim, self.heads
        i64, f32c = _lib.C.c_int64, _lib.C.c_float
        if not speech.is_cuda:
            raise RuntimeError("f5e_tts_b200.ASRModel has no CPU path: pass CUDA tensors")
        B, T, Fdim = speech.shape
        assert Fdim == self.input_dim
        lens = speech_lengths.to(dev).to(torch.int32).flatten()
        if lens.numel() == 1 and B > 1:
            lens = lens.expand(B)
        lens = lens.clamp(max=T).contiguous()
        cm = None if cmvn_applied else self.global_cmvn
        feats = speech.float()
        if cm is not None:  # unfused GlobalCMVN (kaldiFbank.forward(cmvn=...) fuses it)
            feats = (feats - cm.mean.to(dev)) * cm.istd.to(dev)
        feats = feats.contiguous()
        if T < 3:
            return torch.zeros(B, 0, d, device=dev), None
        if not self.use_cuda_graph:
            return self._encode(feats, lens), None
        # one CUDA graph per (batch, frames): the encoder is ~200 small launches, host-launch-bound when issued one by one
        key = (B, T)
        if key not in self._graphs:
            sf, sl = feats.clone(), lens.clone()
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._encode(sf, sl)  # warm-up: position tables, kernel attributes, allocator
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                so = self._encode(sf, sl)
            if len(self._graphs) >= self.max_cached_graphs:  # every graph pins its activations: keep the most recent shapes only
                self._graphs.pop(next(iter(self._graphs)))
            self._graphs[key] = (graph, sf, sl, so)
        graph, sf, sl, so = self._graphs[key]
        sf.copy_(feats)
        sl.copy_(lens)
        graph.replay()
        return so.clone(), None

    def _encode(self, feats: torch.Tensor, lens: torch.Tensor) -> torch.Tensor:
        """Device work of `extract`: no host synchronisation (the padded length comes from T, not from the lengths)."""
        P = self._get_plan()
        lib, dev, d, H = _lib.load(), self.device, self.dim, self.heads
        i64, f32c = _lib.C.c_int64, _lib.C.c_float
        B, T, Fdim = feats.shape
        t_out = torch.where(lens >= 3, (lens - 3) // 2 + 1, torch.zeros_like(lens)).to(torch.int32).contiguous()
        t_max = (T - 3) // 2 + 1  # = (T - 1) // 2: the length the reference returns for a T-frame batch
        n_pad = (t_max + 7) // 8 * 8
        rows = B * n_pad
        st = stream_ptr()
        a0 = torch.empty(rows, P["fo"] * d, device=dev, dtype=torch.bfloat16)
        check(lib.f5e_subsample_conv(ptr(feats), i64(feats.stride(0)), ptr(lens), B, T, Fdim, ptr(P["conv_w"]), ptr(P["conv_b"]), d,
                                     ptr(a0), n_pad, st), "f5e_subsample_conv")
        x = ops.gemm(a0, P["w_out"], bias=P["b_out"])
        del a0
        pos = self._pos_tables(n_pad)
        ld_bias = (n_pad + 127) // 128 * 128
        key_bias = torch.zeros(B * H, ld_bias, device=dev, dtype=torch.float32)
        att = torch.zeros(rows, d, device=dev, dtype=torch.bfloat16)
        conv = torch.empty(rows, d, device=dev, dtype=torch.bfloat16)
        taps = self.conf["cnn_module_kernel"]
        lpad = taps - 1 if self.conf["causal"] else (taps - 1) // 2
        ln = lambda v, wb: ops.norm_mod(v, wb[0], wb[1], plus_one=0.0, eps=1e-5)
        for ly, pl in zip(P["layers"], pos):
            h = ln(x, ly["norm_ff_macaron"])  # encoder_layer.py:206-213
            x = ops.gemm(ops.gemm(h, ly["mw1"], bias=ly["mb1"], act=_lib.ACT_SILU), ly["mw2"], bias=ly["mb2"], residual=x)
            h = ln(x, ly["norm_mha"])  # :216-236
            qkv = ops.gemm(h, ly["wqkv"], bias=ly["bqkv"])
            check(lib.f5e_relpos_prepare(ptr(qkv), qkv.stride(0), ptr(pl), ptr(ly["u"]), ptr(ly["v"]), ptr(key_bias), ld_bias, B,
                                         n_pad, H, st), "f5e_relpos_prepare")
            check(lib.f5e_attention_bias(ptr(qkv), qkv.stride(0), ptr(att), att.stride(0), ptr(t_out), ptr(key_bias), ld_bias, B,
                                         n_pad, H, f32c(1.0 / math.sqrt(d // H)), st), "f5e_attention_bias")
            x = ops.gemm(att, ly["wo"], bias=ly["bo"], residual=x)
            h = ln(x, ly["norm_conv"])  # :243-248
            g = ops.gemm(h, ly["wpw1"], bias=ly["bpw1"])
            check(lib.f5e_glu_dwconv(ptr(g), ptr(ly["wdw"]), ptr(ly["bdw"]), ptr(t_out), ptr(conv), B, n_pad, d, taps, lpad,
                                     ptr(ly["pad"]), ptr(ly["ln_w"]), ptr(ly["ln_b"]), f32c(1e-5), st), "f5e_glu_dwconv")
            x = ops.gemm(conv, ly["wpw2"], bias=ly["bpw2"], residual=x)
            h = ln(x, ly["norm_ff"])  # :254-263
            x = ops.gemm(ops.gemm(h, ly["fw1"], bias=ly["fb1"], act=_lib.ACT_SILU), ly["fw2"], bias=ly["fb2"], residual=x)
            x = ln(x, ly["norm_final"])
        x = ln(x, (P["an_w"], P["an_b"]))  # encoder.py:203-204
        out = ops.gemm(x, P["w_lin"], bias=P["b_lin"], out_f32=True).view(B, n_pad, d)[:, :t_max]  # asr_model.py:241
        valid = torch.arange(t_max, device=dev).unsqueeze(0) < t_out.unsqueeze(1)
        return out * valid.unsqueeze(-1)


def _load_table(x, what):
    """A table given as a tensor / array, or the path the reference loads it from (np.load / pickle, ppg_model.py:82-92)."""
    if isinstance(x, (str, os.PathLike)):
        if str(x).endswith(".npy"):
            import numpy as np
            x = np.load(x)
        else:
            import pickle
            with open(x, "rb") as f:
                x = pickle.load(f)
    if x is None:
        raise ValueError(f"f5e_tts_b200.PPGModelWapper(output_type='map') needs {what} (a tensor / array or the file the "
                         "reference loads; the reference's default paths point outside the tree)")
    return x


class PPGModelWapper(object):
    """ppg/ppg_model.py:59-168.  Pass either the reference's `ppg_model_path` + `ppg_config` (train.yaml + checkpoint) or a
    ready `model`.  output_type "map" (softmax over a phone dictionary, then the phone-centre projection, :113-121) takes its
    two tables as arguments - `global_phn_center` [dict, ppg_dim] and `para_softmax` {'w': [dict, ppg_dim], 'b': [dict]},
    tensors / arrays or the .npy / pickle paths the reference's keyword arguments of the same names hold."""

    def __init__(self, ppg_model_path=None, ppg_config=None, device="cuda", output_type="ppg", map_mix_ratio=1.0,
                 ppg_frame_length=20, mel_f_shift=10, model: ASRModel | None = None, global_phn_center=None, para_softmax=None,
                 global_phn_center_path=None, para_softmax_path=None, **_unused):
        if output_type not in ("ppg", "map"):
            raise ValueError(f"f5e_tts_b200.PPGModelWapper: output_type must be 'ppg' or 'map', not {output_type!r}")
        self._map = None
        if output_type == "map":
            center = torch.as_tensor(_load_table(global_phn_center if global_phn_center is not None else global_phn_center_path,
                                                 "global_phn_center")).float()
            ps = _load_table(para_softmax if para_softmax is not None else para_softmax_path, "para_softmax")
            w, b = torch.as_tensor(ps["w"]).float(), torch.as_tensor(ps["b"]).float()
            V, dim = w.shape
            assert center.shape == (V, dim) and b.shape == (V,), (center.shape, w.shape, b.shape)
            from .vocos import split3_weight
            dev = torch.device(device)
            kpad = (V + 63) // 64 * 64  # K of the second GEMM, padded to whole 64-column blocks
            ct = torch.zeros(dim, kpad)
            ct[:, :V] = center.t()
            self._map = dict(V=V, dim=dim, kpad=kpad, ldl=(V + 3) // 4 * 4, w3=split3_weight(w.to(dev)), b=b.to(dev).contiguous(),
                             c3=split3_weight(ct.to(dev)))
        if model is None:
            import yaml
            with open(ppg_config, "r") as fin:
                conf = yaml.load(fin, Loader=yaml.FullLoader)
            model = ASRModel(conf["input_dim"], **conf.get("encoder_conf", {}))
            model.load_state_dict(torch.load(ppg_model_path, map_location="cpu"))
        self.ppg_model = model.to(device).eval()
        self.device = torch.device(device)
        self.output_type, self.map_mix_ratio = output_type, map_mix_ratio
        self.ppg_frame_length, self.mel_f_shift = ppg_frame_length, mel_f_shift
        self.featCal = kaldiFbank().eval()

    @torch.no_grad()
    def audio_to_mel(self, audio, sr=None, lens=None, *, fuse_cmvn: bool = False):
        """ppg_model.py:147-160: -> (kaldi fbank [B, T, 80], frames).  `lens` (samples) makes the batch ragged; fuse_cmvn applies
        the encoder's GlobalCMVN in the same kernel (then call mel_to_ppg(..., cmvn_applied=True))."""
        if isinstance(audio, str):
            raise NotImplementedError("f5e_tts_b200.PPGModelWapper: pass a waveform tensor (file decoding is outside the hot path)")
        if audio.dim() == 1:
            audio = audio.unsqueeze(0)
        audio = audio.to(self.device)
        if sr is not None and sr != 16000:
            from .pipeline import resample
            audio, new_lens = resample(audio, sr, 16000, lens)
            if lens is not None:
                lens = new_lens.to(self.device)
        feats, feats_len = self.featCal(audio, lens=lens, cmvn=self.ppg_model.global_cmvn if fuse_cmvn else None)
        return feats, feats_len.to(self.device)

    @torch.no_grad()
    def mel_to_ppg(self, mel, mel_lens, *, cmvn_applied: bool = False):
        """ppg_model.py:134-145: fbank features (as audio_to_mel returns them) -> (ppg [B, T', dim] zero beyond true_len, true_len)."""
        ppg, _ = self.ppg_model.extract(mel, mel_lens, stream=False, cmvn_applied=cmvn_applied)
        lens = mel_lens.to(self.device).flatten()
        if lens.numel() == 1 and ppg.shape[0] > 1:
            lens = lens.expand(ppg.shape[0])
        own = torch.where(lens >= 3, (lens - 3) // 2 + 1, torch.zeros_like(lens))
        true_len = torch.minimum(ppg_true_len(lens, ppg.shape[1], self.ppg_frame_length, self.mel_f_shift), own)
        ppg = self.ppg_to_target(ppg, true_len)
        return ppg, true_len

    @torch.no_grad()
    def ppg_to_target(self, ppg, true_len):
        """ppg_model.py:108-127: output_type "ppg" masks the padded frames; "map" first projects every frame onto the phone
        centres: softmax(ppg @ w^T + b) @ global_phn_center, mixed with the raw PPG by map_mix_ratio.  Both GEMMs are the
        fp32-accurate bf16x3 tcgen05 GEMM, the softmax one warp per frame."""
        mask = (~make_pad_mask(true_len, ppg.shape[1]))[:, :, None]
        if self._map is None:
            return ppg * mask
        P = self._map
        B, T, dim = ppg.shape
        assert dim == P["dim"], (dim, P["dim"])
        lib, st, M = _lib.load(), stream_ptr(), B * T
        x = ppg.reshape(M, dim).float().contiguous()
        a3 = torch.empty(M, 3 * dim, device=x.device, dtype=torch.bfloat16)
        check(lib.f5e_split3(ptr(x), dim, M, dim, ptr(a3), st), "f5e_split3")
        logits = ops.gemm(a3, P["w3"], bias=P["b"], out=torch.empty(M, P["ldl"], device=x.device), n=P["V"])
        p3 = torch.empty(M, 3 * P["kpad"], device=x.device, dtype=torch.bfloat16)
        check(lib.f5e_softmax_split3(ptr(logits), P["ldl"], M, P["V"], P["kpad"], ptr(p3), st), "f5e_softmax_split3")
        mapped = ops.gemm(p3, P["c3"], out_f32=True).view(B, T, dim)
        if self.map_mix_ratio != 1.0:
            mapped = ppg * (1 - self.map_mix_ratio) + mapped * self.map_mix_ratio
        return mapped * mask

    @torch.no_grad()
    def audio_to_ppg(self, audio, sr=None, lens=None):
        feats, feats_len = self.audio_to_mel(audio, sr, lens, fuse_cmvn=True)
        return self.mel_to_ppg(feats, feats_len, cmvn_applied=True)
