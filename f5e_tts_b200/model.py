"""Drop-in mirrors of the reference's inference objects, backed by libf5e_b200.so.

    DiT      <-> f5_tts.model.backbones.dit.DiT        (constructor kwargs, .sample, .clear_cache, state-dict keys)
    CFM      <-> f5_tts.model.cfm.CFM                  (.sample -> (out, trajectory))
    MelSpec  <-> f5_tts.model.modules.MelSpec          (see frontend.py)

torch supplies parameter storage (so `load_state_dict` / `load_checkpoint` of the reference work unchanged),
device memory and the current stream.  All arithmetic on the hot path runs in the hand-written sm_100a
kernels reached through ctypes; there is no eager fallback - on a CPU tensor or without the library the
calls raise.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import threading

import torch
import torch.nn.functional as F
from torch import nn

from . import _lib, ops
from ._lib import check, ptr, stream_ptr

_SOLVE_LOCK = threading.RLock()


# ---- ctypes mirrors of include/f5e_b200.h ------------------------------------------------------------
class _Layer(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("w_qkv", "b_qkv", "w_out", "b_out", "w_ff1", "b_ff1", "w_ff2", "b_ff2",
                                          "wf_qkv", "c_qkv", "d_qkv", "wf_ff1", "c_ff1", "d_ff1", "q_norm_w", "k_norm_w")]


class _FoldJob(C.Structure):  # f5e_fold_job
    _fields_ = [("w", C.c_void_p), ("bias", C.c_void_p), ("wf", C.c_void_p), ("c", C.c_void_p), ("d", C.c_void_p),
                ("N", C.c_int32), ("row0", C.c_int32), ("scale_off", C.c_int32), ("shift_off", C.c_int32)]


class _Model(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("dim", "depth", "heads", "ff_dim", "in_dim", "mel_dim", "rope_heads",
                                         "conv_groups")] + [
        ("w_in", C.c_void_p), ("b_in", C.c_void_p), ("w_conv1", C.c_void_p), ("b_conv1", C.c_void_p),
        ("w_conv2", C.c_void_p), ("b_conv2", C.c_void_p), ("layers", C.POINTER(_Layer)),
        ("w_proj", C.c_void_p), ("b_proj", C.c_void_p), ("rope_inv_freq", C.c_void_p), ("fold_jobs", C.c_void_p), ("fold_n_jobs", C.c_int32), ("fold_rows", C.c_int32),
        ("w_long_skip", C.c_void_p), ("isolate_samples", C.c_int32)]


class _Buffers(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("x_cat", "h0", "h1", "x", "xn", "qkv", "attn", "ff", "pred", "stats", "stats_a",
                                          "stats_part")]


class _Packing(C.Structure):  # f5e_dit_packing
    _fields_ = [("row_map", C.c_void_p), ("row_info", C.c_void_p), ("row_start", C.c_void_p), ("rows_packed", C.c_int32),
                ("rows_total", C.c_int32)]


def make_packing(klen: torch.Tensor, n_pad: int, keep_padded: bool = False):
    """Row packing of `samples` sequences with klen[s] valid frames each (include/f5e_b200.h: f5e_dit_packing).
    keep_padded: the padded frames follow the valid ones (the exact, "split" form) instead of being dropped.
    Integer index plumbing in torch; returns (struct, tensors to keep alive)."""
    klen = klen.to(torch.int64)
    S = klen.numel()
    frame = torch.arange(n_pad, device=klen.device)
    valid = frame[None, :] < klen[:, None]                                   # [S, n_pad]
    sample = torch.arange(S, device=klen.device)[:, None].expand(S, n_pad)
    rm, ri = sample * n_pad + frame[None, :], (sample << 16) | frame[None, :]
    row_map, row_info = rm[valid], ri[valid]
    rows_packed = int(row_map.numel())
    if keep_padded:
        row_map, row_info = torch.cat([row_map, rm[~valid]]), torch.cat([row_info, ri[~valid]])
    row_map, row_info = row_map.to(torch.int32).contiguous(), row_info.to(torch.int32).contiguous()
    row_start = (torch.cumsum(klen, 0) - klen).to(torch.int32).contiguous()
    pk = _Packing(row_map.data_ptr(), row_info.data_ptr(), row_start.data_ptr(), rows_packed,
                  int(row_map.numel()) if keep_padded else 0)
    return pk, (row_map, row_info, row_start)


# ---- parameter containers with the reference's state-dict names ---------------------------------------
def _seq(mods: dict) -> nn.ModuleDict:
    """nn.Sequential-compatible naming with gaps ("0", "2", ...)."""
    return nn.ModuleDict({str(k): v for k, v in mods.items()})


class _GRN(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.gamma = nn.Parameter(torch.zeros(1, 1, dim))
        self.beta = nn.Parameter(torch.zeros(1, 1, dim))


class _ConvNeXtV2Block(nn.Module):  # parameters of model/modules.py:241-259
    def __init__(self, dim, inter):
        super().__init__()
        self.dwconv = nn.Conv1d(dim, dim, 7, padding=3, groups=dim)
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.pwconv1 = nn.Linear(dim, inter)
        self.grn = _GRN(inter)
        self.pwconv2 = nn.Linear(inter, dim)


class _TextEmbedding(nn.Module):  # backbones/dit.py:37-52
    def __init__(self, n, dim, conv_layers):
        super().__init__()
        self.text_embed = nn.Embedding(n + 1, dim)
        if conv_layers > 0:
            self.text_blocks = nn.ModuleList([_ConvNeXtV2Block(dim, dim * 2) for _ in range(conv_layers)])


class _Holder(nn.Module):
    def __init__(self, **mods):
        super().__init__()
        for k, v in mods.items():
            setattr(self, k, v)


class _RMSNormWeight(nn.Module):  # parameter container of model/modules.py:275-280 (the arithmetic is qk_norm_rope_kernel)
    def __init__(self, dim):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))


class _Block(nn.Module):  # model/modules.py:610-625
    def __init__(self, dim, heads, dim_head, ff_mult, qk_norm=None):
        super().__init__()
        inner = heads * dim_head
        self.attn_norm = _Holder(linear=nn.Linear(dim, dim * 6))
        self.attn = _Holder(to_q=nn.Linear(dim, inner), to_k=nn.Linear(dim, inner), to_v=nn.Linear(dim, inner),
                            to_out=_seq({0: nn.Linear(inner, dim)}))
        if qk_norm == "rms_norm":  # modules.py:394-396
            self.attn.q_norm, self.attn.k_norm = _RMSNormWeight(dim_head), _RMSNormWeight(dim_head)
        self.ff = _Holder(ff=_seq({0: _seq({0: nn.Linear(dim, dim * ff_mult)}), 2: nn.Linear(dim * ff_mult, dim)}))


def _sinus_time_embed(t: torch.Tensor, dim=256, scale=1000) -> torch.Tensor:
    """model/modules.py:149-161 with the same dtype promotion (`scale * t` in t's dtype, then fp32).  33 scalars of
    host-side schedule glue; kept in torch so the rounding matches the reference op for op."""
    half = dim // 2
    e = math.log(10000) / (half - 1)
    e = torch.exp(torch.arange(half, device=t.device).float() * -e)
    e = scale * t.unsqueeze(1) * e.unsqueeze(0)
    return torch.cat((e.sin(), e.cos()), dim=-1)


class DiT(nn.Module):
    """Mirror of reference backbones/dit.py:183-271 (constructor) and :417-472 (`sample`)."""

    def __init__(self, *, dim, depth=8, heads=8, dim_head=64, dropout=0.1, ff_mult=4, mel_dim=100, text_num_embeds=256,
                 text_dim=None, text_mask_padding=True, qk_norm=None, conv_layers=0, pe_attn_head=None,
                 long_skip_connection=False, checkpoint_activations=False, ppg_config=dict(use_ppg=False),
                 cb_config=dict(use_codebook=False)):
        super().__init__()
        if dim_head != 64:
            raise NotImplementedError("f5e_tts_b200: the attention kernel is specialised for dim_head=64")
        if qk_norm not in (None, "rms_norm"):
            raise ValueError(f"Unimplemented qk_norm: {qk_norm}")  # as the reference (modules.py:397-398)
        if dim % 256 or (dim // 16) % 16 or dim // 16 > 64:
            raise NotImplementedError(f"f5e_tts_b200: dim={dim} unsupported (need dim % 256 == 0 and dim/16 in 16..64)")
        text_dim = mel_dim if text_dim is None else text_dim
        self.dim, self.depth, self.heads, self.dim_head, self.ff_mult = dim, depth, heads, dim_head, ff_mult
        self.mel_dim, self.text_dim, self.text_num_embeds = mel_dim, text_dim, text_num_embeds
        self.text_mask_padding, self.conv_layers, self.pe_attn_head = text_mask_padding, conv_layers, pe_attn_head
        self.use_ppg = ppg_config["use_ppg"]
        self.use_codebook = cb_config["use_codebook"]

        self.time_embed = _Holder(time_mlp=_seq({0: nn.Linear(256, dim), 2: nn.Linear(dim, dim)}))
        self.text_embed = _TextEmbedding(text_num_embeds, text_dim, conv_layers)
        in_dim = mel_dim * 2 + text_dim * (2 if self.use_ppg else 1)
        self.ppg_transformer = False
        if self.use_ppg:
            P = self.ppg_dim = ppg_config["ppg_dim"]
            self.ppg_transformer = bool(ppg_config.get("use_transformer", False))
            if self.ppg_transformer:  # dit.py:105-119; torch's modules are used as parameter containers only (same keys)
                tc = ppg_config["transformer_config"]
                if P % tc["nhead"] or P // tc["nhead"] != 64 or P % 256:
                    raise NotImplementedError("f5e_tts_b200: the transformer PPG embedder needs ppg_dim % 256 == 0 and "
                                              "ppg_dim / nhead == 64 (the attention kernel's head width)")
                layer = nn.TransformerEncoderLayer(d_model=P, nhead=tc["nhead"], dim_feedforward=tc["dim_feedforward"],
                                                   dropout=tc["dropout"], activation=F.gelu, batch_first=True)
                self.ppg_embed = _Holder(ppg_proj=_seq({0: nn.TransformerEncoder(layer, num_layers=tc["num_layers"]),
                                                        1: nn.Linear(P, text_dim)}))
            else:
                self.ppg_embed = _Holder(ppg_proj=_seq({0: nn.Linear(P, P), 2: nn.Conv1d(P, P, 5, padding="same"),
                                                        3: nn.BatchNorm1d(P), 6: nn.Conv1d(P, P, 5, padding="same"),
                                                        7: nn.BatchNorm1d(P), 10: nn.Conv1d(P, P, 5, padding="same"),
                                                        11: nn.BatchNorm1d(P), 15: nn.Linear(P, text_dim)}))
        if self.use_codebook:
            from .quantizer import GumbelVectorQuantizer
            self.quantizer = GumbelVectorQuantizer(
                dim=text_dim, num_vars=cb_config["num_vars"],
                temp=(cb_config["temp_start"], cb_config["temp_stop"], cb_config["temp_decay"]),
                groups=cb_config["groups"], combine_groups=cb_config["combine_groups"], vq_dim=text_dim,
                time_first=True, weight_proj_depth=cb_config["weight_proj_depth"],
                weight_proj_factor=cb_config["weight_proj_factor"])
        self.in_dim = in_dim
        self.input_embed = _Holder(proj=nn.Linear(in_dim, dim), conv_pos_embed=_Holder(conv1d=_seq({
            0: nn.Conv1d(dim, dim, 31, groups=16, padding=15), 2: nn.Conv1d(dim, dim, 31, groups=16, padding=15)})))
        self.rotary_embed = _Holder()
        self.rotary_embed.register_buffer("inv_freq", 1.0 / (10000 ** (torch.arange(0, dim_head, 2).float() / dim_head)))
        self.qk_norm = qk_norm
        self.transformer_blocks = nn.ModuleList([_Block(dim, heads, dim_head, ff_mult, qk_norm) for _ in range(depth)])
        self.long_skip_connection = nn.Linear(dim * 2, dim, bias=False) if long_skip_connection else None  # dit.py:264
        self.norm_out = _Holder(linear=nn.Linear(dim, dim * 2))
        self.proj_out = nn.Linear(dim, mel_dim)
        self.text_cond = self.text_uncond = None
        self._plan = None
        self._ws = {}
        # Euler loop only: fold the adaLN LayerNorms into the QKV / FF1 GEMMs (no normalised copy of the activations)
        self.fold_layernorm = os.environ.get("F5E_FOLD_LN", "1") != "0"
        self.fold_min_rows = int(os.environ.get("F5E_FOLD_MIN_ROWS", "6144"))
        self._gbuf = {}  # persistent buffers of the cached-CUDA-graph path, per solve shape
        self.initialize_weights()

    def initialize_weights(self):  # backbones/dit.py:273-283
        for blk in self.transformer_blocks:
            nn.init.constant_(blk.attn_norm.linear.weight, 0)
            nn.init.constant_(blk.attn_norm.linear.bias, 0)
        for m in (self.norm_out.linear, self.proj_out):
            nn.init.constant_(m.weight, 0)
            nn.init.constant_(m.bias, 0)

    def clear_cache(self):  # backbones/dit.py:293-294
        self.text_cond = self.text_uncond = None

    # Any parameter change invalidates the packed plan: `_apply` (.to / .half / .cuda), `load_state_dict` on this module
    # OR on a parent (the reference's load_checkpoint loads into the CFM: infer/utils_infer.py:196-234, which reaches the
    # children through _load_from_state_dict only), and in-place updates (EMA copy_, optimizer steps), which bump the
    # parameters' version counters.
    def _drop_plan(self):
        if getattr(self, "_gbuf", None):  # cached graphs refer to the plan's weights and to these buffers
            _lib.load().f5e_graph_cache_clear()
        self._plan, self._ws, self._gbuf = None, {}, {}

    def _apply(self, fn, *a, **k):
        self._drop_plan()
        return super()._apply(fn, *a, **k)

    def _load_from_state_dict(self, *a, **k):
        self._drop_plan()
        return super()._load_from_state_dict(*a, **k)

    def _param_versions(self):
        return tuple(p._version for p in self.parameters()) + tuple(b._version for b in self.buffers())

    def forward(self, *a, **k):
        raise NotImplementedError("f5e_tts_b200.DiT is inference-only: use .sample (training stays with the reference)")

    # ---- packed weights ("plan") ------------------------------------------------------------------
    def _get_plan(self):
        if self._plan is not None:
            if self._plan.versions == self._param_versions():
                return self._plan
            self._drop_plan()
        dev = self.proj_out.weight.device
        if dev.type != "cuda":
            raise RuntimeError("f5e_tts_b200.DiT has no CPU path: move the module to a B200 (cuda) device")
        _lib.require_device()
        bf = lambda t: t.detach().to(dev, torch.bfloat16).contiguous()  # noqa: E731
        f32 = lambda t: t.detach().to(dev, torch.float32).contiguous()  # noqa: E731
        P = type("Plan", (), {})()
        keep = P.keep = []

        def K(t):
            keep.append(t)
            return t.data_ptr()

        D = self.dim
        layers = (_Layer * self.depth)()
        mod_w, mod_b = [], []
        for l, blk in enumerate(self.transformer_blocks):
            a = blk.attn
            layers[l].w_qkv = K(bf(torch.cat([a.to_q.weight, a.to_k.weight, a.to_v.weight], 0)))
            layers[l].b_qkv = K(f32(torch.cat([a.to_q.bias, a.to_k.bias, a.to_v.bias], 0)))
            layers[l].w_out, layers[l].b_out = K(bf(a.to_out["0"].weight)), K(f32(a.to_out["0"].bias))
            layers[l].w_ff1, layers[l].b_ff1 = K(bf(blk.ff.ff["0"]["0"].weight)), K(f32(blk.ff.ff["0"]["0"].bias))
            layers[l].w_ff2, layers[l].b_ff2 = K(bf(blk.ff.ff["2"].weight)), K(f32(blk.ff.ff["2"].bias))
            if self.qk_norm == "rms_norm":
                layers[l].q_norm_w, layers[l].k_norm_w = K(f32(a.q_norm.weight)), K(f32(a.k_norm.weight))
            mod_w.append(blk.attn_norm.linear.weight)
            mod_b.append(blk.attn_norm.linear.bias)
        # folded-LayerNorm path of the Euler loop (include/f5e_b200.h: f5e_fold_job): per-step rewritten QKV / FF1 weights
        jobs = (_FoldJob * (2 * self.depth))()
        row0 = 0
        for l in range(self.depth):
            for which, (wk, bk, n_out, sc, sh) in enumerate((("w_qkv", "b_qkv", 3 * self.heads * 64, 1, 0),
                                                             ("w_ff1", "b_ff1", D * self.ff_mult, 4, 3))):
                wf = torch.zeros(n_out, D + 64, device=dev, dtype=torch.bfloat16)
                cv, dv = torch.empty(n_out, device=dev), torch.empty(n_out, device=dev)
                tag = "qkv" if which == 0 else "ff1"
                setattr(layers[l], "wf_" + tag, K(wf))
                setattr(layers[l], "c_" + tag, K(cv))
                setattr(layers[l], "d_" + tag, K(dv))
                j = jobs[2 * l + which]
                j.w, j.bias, j.wf, j.c, j.d = getattr(layers[l], wk), getattr(layers[l], bk), wf.data_ptr(), cv.data_ptr(), dv.data_ptr()
                j.N, j.row0, j.scale_off, j.shift_off = n_out, row0, (l * 6 + sc) * D, (l * 6 + sh) * D
                row0 += n_out
        P.fold_jobs = torch.frombuffer(bytearray(bytes(jobs)), dtype=torch.uint8).to(dev)
        P.fold_n_jobs, P.fold_rows = 2 * self.depth, row0
        mod_w.append(self.norm_out.linear.weight)
        mod_b.append(self.norm_out.linear.bias)
        P.w_mod, P.b_mod = bf(torch.cat(mod_w, 0)), f32(torch.cat(mod_b, 0))
        P.mod_size = P.w_mod.shape[0]
        tm = self.time_embed.time_mlp
        P.w_t0, P.b_t0, P.w_t2, P.b_t2 = bf(tm["0"].weight), f32(tm["0"].bias), bf(tm["2"].weight), f32(tm["2"].bias)
        m = _Model()
        m.dim, m.depth, m.heads, m.ff_dim, m.in_dim, m.mel_dim = D, self.depth, self.heads, D * self.ff_mult, self.in_dim, self.mel_dim
        m.rope_heads = self.heads if self.pe_attn_head is None else self.pe_attn_head
        m.conv_groups = 16
        ie = self.input_embed
        m.w_in, m.b_in = K(bf(ie.proj.weight)), K(f32(ie.proj.bias))
        c = ie.conv_pos_embed.conv1d
        m.w_conv1, m.b_conv1 = K(ops.pack_conv_pos_weight(c["0"].weight.detach().to(dev))), K(f32(c["0"].bias))
        m.w_conv2, m.b_conv2 = K(ops.pack_conv_pos_weight(c["2"].weight.detach().to(dev))), K(f32(c["2"].bias))
        m.layers = C.cast(layers, C.POINTER(_Layer))
        m.w_proj, m.b_proj = K(bf(self.proj_out.weight)), K(f32(self.proj_out.bias))
        if self.long_skip_connection is not None:
            m.w_long_skip = K(bf(self.long_skip_connection.weight))
        m.rope_inv_freq = K(f32(self.rotary_embed.inv_freq))  # angles are evaluated in the QKV epilogue
        P.layers, P.model = layers, m
        # text embedder
        te = self.text_embed
        P.table = bf(te.text_embed.weight)
        P.text_blocks = []
        if self.conv_layers > 0:
            from .tables import text_pos_table
            P.pos = text_pos_table(self.text_dim).to(dev)
            for b in te.text_blocks:
                P.text_blocks.append(dict(dw_w=f32(b.dwconv.weight.reshape(self.text_dim, 7)), dw_b=f32(b.dwconv.bias),
                                          ln_w=f32(b.norm.weight), ln_b=f32(b.norm.bias),
                                          w1=bf(b.pwconv1.weight), b1=f32(b.pwconv1.bias),
                                          gamma=f32(b.grn.gamma.reshape(-1)), beta=f32(b.grn.beta.reshape(-1)),
                                          w2=bf(b.pwconv2.weight), b2=f32(b.pwconv2.bias)))
        else:
            P.pos = None
        if self.use_ppg and self.ppg_transformer:
            pp = self.ppg_embed.ppg_proj
            P.ppg = dict(w1=bf(pp["1"].weight), b1=f32(pp["1"].bias), heads=pp["0"].layers[0].self_attn.num_heads, layers=[])
            for ly in pp["0"].layers:
                sa = ly.self_attn
                P.ppg["layers"].append(dict(
                    w_in=bf(sa.in_proj_weight), b_in=f32(sa.in_proj_bias), w_out=bf(sa.out_proj.weight), b_out=f32(sa.out_proj.bias),
                    w_l1=bf(ly.linear1.weight), b_l1=f32(ly.linear1.bias), w_l2=bf(ly.linear2.weight), b_l2=f32(ly.linear2.bias),
                    n1w=f32(ly.norm1.weight), n1b=f32(ly.norm1.bias), n2w=f32(ly.norm2.weight), n2b=f32(ly.norm2.bias),
                    eps=ly.norm1.eps))
        elif self.use_ppg:
            pp = self.ppg_embed.ppg_proj
            P.ppg = dict(w0=bf(pp["0"].weight), b0=f32(pp["0"].bias), w15=bf(pp["15"].weight), b15=f32(pp["15"].bias), convs=[])
            for ci, bi in (("2", "3"), ("6", "7"), ("10", "11")):
                conv, bn = pp[ci], pp[bi]
                s = (bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)).to(dev)
                w = conv.weight.detach().float().to(dev) * s[:, None, None]                # fold eval-mode BatchNorm
                b = (conv.bias.detach().float().to(dev) - bn.running_mean.detach().float().to(dev)) * s + bn.bias.detach().float().to(dev)
                P.ppg["convs"].append((bf(w.permute(0, 2, 1).reshape(w.shape[0], -1)), f32(b)))  # [co, tap*C + ci]
        P.versions = self._param_versions()
        pd = self.proj_out.weight.dtype
        if pd != torch.bfloat16 and not getattr(DiT, "_warned_dtype", False):
            # the reference's load_checkpoint casts to fp16 on the GPU (infer/utils_infer.py:185-194); see DESIGN.md 2.
            import warnings
            warnings.warn(f"f5e_tts_b200.DiT: parameters are {pd}; the sm_100a kernels compute in bf16 with fp32 accumulation "
                          "(weights are packed to bf16 once), the ODE state / noise / time grid keep the module dtype",
                          stacklevel=3)
            DiT._warned_dtype = True
        self._plan = P
        return P

    def _workspace(self, samples, n_pad):
        key = (samples, n_pad)
        if key not in self._ws:
            dev, rows, D, inner = self._get_plan().w_mod.device, samples * n_pad, self.dim, self.heads * 64
            mk = lambda c, dt=torch.bfloat16: torch.empty(rows, c, device=dev, dtype=dt)  # noqa: E731
            t = dict(x_cat=torch.zeros(rows, self.in_dim, device=dev, dtype=torch.bfloat16), h0=mk(D), h1=mk(D), x=mk(D),
                     xn=mk(D), qkv=mk(3 * inner), attn=torch.zeros(rows, inner, device=dev, dtype=torch.bfloat16),
                     ff=mk(D * self.ff_mult), pred=mk(self.mel_dim, torch.float32), stats=mk(1, torch.float32),
                     stats_a=torch.zeros(rows, 64, device=dev, dtype=torch.bfloat16),
                     stats_part=mk(max(1, D // 128) * 2, torch.float32))
            b = _Buffers(**{k: v.data_ptr() for k, v in t.items()})
            if self._gbuf:  # cached graphs refer to the workspace being replaced
                _lib.load().f5e_graph_cache_clear()
                self._gbuf = {}
            self._ws = {key: (t, b)}  # keep only the latest shape
        return self._ws[key]

    # ---- once-per-solve conditioning ----------------------------------------------------------------
    def _text_embed(self, text, batch, seq_len, drop_text):
        """TextEmbedding.forward (backbones/dit.py:54-87) -> bf16 [B, N, text_dim]."""
        P = self._get_plan()
        if text is None:
            text = torch.full((batch, 1), -1, dtype=torch.long, device=P.table.device)
            mask_padding = False  # reference: no mask is built when text is None
        else:
            mask_padding = self.text_mask_padding
        text = text.to(P.table.device)
        x = ops.text_gather(text, P.table, P.pos, seq_len, drop=drop_text, mask_padding=mask_padding and self.conv_layers > 0,
                            add_pos=self.conv_layers > 0)
        if self.conv_layers == 0:
            return x
        row_len = None
        if mask_padding:
            valid = (text[:, :seq_len] != -1)
            tl = valid.sum(-1)
            if not bool((valid == (torch.arange(valid.shape[1], device=valid.device)[None] < tl[:, None])).all()):
                raise NotImplementedError("f5e_tts_b200: text padding (-1) must be trailing")
            row_len = tl.to(torch.int32)
        B, N, TD = x.shape
        for blk in P.text_blocks:  # ConvNeXtV2Block (model/modules.py:261-269)
            h = ops.dwconv7(x, blk["dw_w"], blk["dw_b"])
            h = ops.norm_mod(h, blk["ln_w"], blk["ln_b"], plus_one=0.0, eps=1e-6)
            h = ops.gemm(h.view(B * N, TD), blk["w1"], bias=blk["b1"], act=_lib.ACT_GELU_ERF)
            h = ops.grn(h.view(B, N, -1), blk["gamma"], blk["beta"])
            # residual + pwconv2; rows beyond the text are zero in `x` and zeroed in the branch -> masked_fill(0)
            x = ops.gemm(h.view(B * N, -1), blk["w2"], bias=blk["b2"], residual=x.view(B * N, TD), row_len=row_len,
                         n_pad=N).view(B, N, TD)
        return x

    def _ppg_embed(self, ppg, seq_len, drop_ppg, batch):
        """PPGEmbedding.forward (backbones/dit.py:140-153), BatchNorm folded into the k=5 convs."""
        P = self._get_plan()
        dev, Pd = P.table.device, self.ppg_dim
        x = torch.zeros(batch, seq_len, Pd, device=dev, dtype=torch.bfloat16)
        if ppg is not None and not drop_ppg:
            n = min(ppg.shape[1], seq_len)
            ops.pack_cols(x.view(batch * seq_len, Pd).view(batch, seq_len * Pd)[:, : n * Pd], 0,
                          ppg[:, :n].reshape(batch, n * Pd).to(dev))
        rows = batch * seq_len
        if self.ppg_transformer:  # post-norm nn.TransformerEncoderLayer, exact GELU, no padding mask (dit.py:105-119)
            h = x.view(rows, Pd)
            for ly in P.ppg["layers"]:
                qkv = ops.gemm(h, ly["w_in"], bias=ly["b_in"])
                a = ops.attention(qkv, P.ppg["heads"], n_pad=seq_len)
                h = ops.gemm(a, ly["w_out"], bias=ly["b_out"], residual=h)
                h = ops.norm_mod(h, ly["n1w"], ly["n1b"], plus_one=0.0, eps=ly["eps"])
                f = ops.gemm(h, ly["w_l1"], bias=ly["b_l1"], act=_lib.ACT_GELU_ERF)
                h = ops.gemm(f, ly["w_l2"], bias=ly["b_l2"], residual=h)
                h = ops.norm_mod(h, ly["n2w"], ly["n2b"], plus_one=0.0, eps=ly["eps"])
            return ops.gemm(h, P.ppg["w1"], bias=P.ppg["b1"]).view(batch, seq_len, self.text_dim)
        h = ops.gemm(x.view(rows, Pd), P.ppg["w0"], bias=P.ppg["b0"])
        for w, b in P.ppg["convs"]:
            u = ops.unfold_frames(h.view(batch, seq_len, Pd), 5)
            h = ops.gemm(u.view(rows, 5 * Pd), w, bias=b, act=_lib.ACT_RELU)
        return ops.gemm(h, P.ppg["w15"], bias=P.ppg["b15"]).view(batch, seq_len, self.text_dim)

    def _modulation(self, times: torch.Tensor) -> torch.Tensor:
        """All time-dependent vectors for the given time values: TimestepEmbedding (model/modules.py:721-731),
        then every AdaLayerNorm / AdaLayerNorm_Final linear on SiLU(t) (:310, :332) in ONE GEMM -> fp32 [S, mod_size]."""
        P = self._get_plan()
        th = _sinus_time_embed(times.to(P.w_mod.device)).to(times.dtype).to(torch.bfloat16).contiguous()
        h = ops.gemm(th, P.w_t0, bias=P.b_t0, act=_lib.ACT_SILU)
        st = ops.gemm(h, P.w_t2, bias=P.b_t2, act=_lib.ACT_SILU)  # SiLU(t_emb): the only way t_emb is consumed
        return ops.gemm(st, P.w_mod, bias=P.b_mod, out_f32=True)

    # ---- one flow evaluation ----------------------------------------------------------------------
    @torch.no_grad()
    def sample(self, x, cond, text, ppg, time, drop_audio_cond, drop_text, drop_ppg, mask=None):
        """backbones/dit.py:417-472."""
        with _SOLVE_LOCK:  # one plan / workspace per model: a direct call must not interleave with a running CFM solve
            return self._sample_locked(x, cond, text, ppg, time, drop_audio_cond, drop_text, drop_ppg, mask)

    def _sample_locked(self, x, cond, text, ppg, time, drop_audio_cond, drop_text, drop_ppg, mask):
        P = self._get_plan()
        B, N = x.shape[0], x.shape[1]
        dev = P.table.device
        if time.ndim == 0:
            time = time.repeat(B)
        mod = self._modulation(time)
        if drop_text:
            if self.text_uncond is None:
                self.text_uncond = self._text_embed(text, B, N, True)
            te = self.text_uncond
        else:
            if self.text_cond is None:
                self.text_cond = self._text_embed(text, B, N, False)
            te = self.text_cond
        tens, buf = self._workspace(B, N)
        xc, md, td = tens["x_cat"], self.mel_dim, self.text_dim
        ops.pack_cols(xc, 0, x.reshape(B * N, md).to(dev))
        ops.pack_cols(xc, md, None if drop_audio_cond else cond.reshape(B * N, md).to(dev), md)
        ops.pack_cols(xc, 2 * md, te.view(B * N, td))
        if self.use_ppg:
            ops.pack_cols(xc, 2 * md + td, self._ppg_embed(ppg, N, drop_ppg, B).view(B * N, td))
        klen = None
        if mask is not None:
            klen = mask.sum(-1).to(torch.int32).to(dev)
        P.model.isolate_samples = 0
        check(_lib.load().f5e_dit_forward(C.byref(P.model), C.byref(buf), ptr(mod), P.mod_size, ptr(klen), B, N, None,
                                          stream_ptr()), "f5e_dit_forward")
        return tens["pred"].view(B, N, md).to(x.dtype)


# =======================================================================================================
# CFM: the sampler
# =======================================================================================================
def lens_to_mask(t, length=None):  # model/utils.py:41-46
    if length is None:
        length = int(t.amax())
    return torch.arange(length, device=t.device)[None, :] < t[:, None]


def list_str_to_idx(text, vocab_char_map, padding_value=-1):  # model/utils.py:87-100
    toks = [torch.tensor([vocab_char_map.get(c, 0) for c in t]) for t in text]
    return torch.nn.utils.rnn.pad_sequence(toks, padding_value=padding_value, batch_first=True)


def list_str_to_tensor(text, padding_value=-1):  # model/utils.py:80-84
    toks = [torch.tensor([*bytes(t, "UTF-8")]) for t in text]
    return torch.nn.utils.rnn.pad_sequence(toks, padding_value=padding_value, batch_first=True)


class CFM(nn.Module):
    """Mirror of reference model/cfm.py:34-91 (constructor / attributes) and :348-482 (`sample`)."""

    def __init__(self, transformer, sigma=0.0, odeint_kwargs=dict(method="euler"), audio_drop_prob=0.3,
                 cond_drop_prob=0.2, num_channels=None, mel_spec_module=None, mel_spec_kwargs=dict(),
                 frac_lengths_mask=(0.7, 1.0), vocab_char_map=None, ppg_config=dict(use_ppg=False),
                 cb_config=dict(use_codebook=False)):
        super().__init__()
        from .frontend import MelSpec
        self.frac_lengths_mask = frac_lengths_mask
        self.mel_spec = mel_spec_module if mel_spec_module is not None else MelSpec(**mel_spec_kwargs)
        self.num_channels = num_channels if num_channels is not None else self.mel_spec.n_mel_channels
        self.audio_drop_prob, self.cond_drop_prob = audio_drop_prob, cond_drop_prob
        self.transformer = transformer
        self.dim = transformer.dim
        self.sigma = sigma
        self.odeint_kwargs = odeint_kwargs
        self.vocab_char_map = vocab_char_map
        self.use_ppg = ppg_config["use_ppg"]
        self.use_cross_mask = ppg_config.get("use_cross_mask", False)
        self.use_codebook = cb_config["use_codebook"]
        self.use_align_loss = cb_config.get("use_align_loss", False)
        # knobs of the B200 path (not in the reference)
        self.use_cuda_graph = False       # capture the whole Euler loop in one CUDA graph launch
        self.store_trajectory = True      # False: return only [y0, y1] instead of steps+1 states
        self.emulate_lowp_state = False   # True: round the ODE state to bf16 every step like a bf16 reference run
        # batch > 1: the DiT blocks run on a permuted row order with the valid frames first (f5e_dit_packing).
        # None (default) = always exact.  The reference's padded frames evolve with every Euler step and leak through the two
        #   unmasked k=31 ConvPositionEmbedding convs into the last <= 30 valid frames of every shorter sample,
        #   so they are kept: they follow the valid rows and skip only the attention sub-block, whose result the reference
        #   masks to zero on them (bit-identical to the padded layout, no QKV / out-projection FLOPs on padding).  Where no
        #   padded frame can reach a valid one (`isolate_samples`, or equal lengths) they are dropped altogether.
        # True = always drop them (throughput knob; their state is then frozen at y0 = 0, the deviation is confined to those
        #   tail frames and measured in tests/test_sampler_options_gpu.py::test_packed_rows_tail_deviation).
        # False = the padded layout everywhere.
        self.pack_valid_rows = None
        # True: a sample's result does not depend on what it is batched with - padded frames are zeroed at the inputs of
        # the ConvPositionEmbedding convs and the text / PPG embedders run per sample at the sample's own length, so a
        # padded batch reproduces the batch-1 solves of infer/utils_infer.py:474-482 (pipeline.infer_batch_process).
        # False (default) = the reference's batched behaviour, padding leaks included.
        self.isolate_samples = False

    @property
    def device(self):
        return next(self.parameters()).device

    def forward(self, *a, **k):
        raise NotImplementedError("f5e_tts_b200.CFM is inference-only: use .sample")

    @torch.no_grad()
    def sample(self, cond, text, ppg=None, duration=None, *, lens=None, steps=32, cfg_strength=1.0,
               sway_sampling_coef=None, seed=None, max_duration=4096, vocoder=None, no_ref_audio=False,
               duplicate_test=False, t_inter=0.1, edit_mask=None):
        """Same contract as reference model/cfm.py:348-482; the Euler/CFG loop runs in libf5e_b200 on a doubled batch
        (conditional half, then the all-dropped half)."""
        if cfg_strength >= 1e-5:
            branches, g = [(False, False, False), (True, True, True)], _lib.Guidance(2, float(cfg_strength), 0.0)
        else:
            branches, g = [(False, False, False)], _lib.Guidance(1, 0.0, 0.0)
        return self._solve(cond, text, ppg, duration, branches, g, lens=lens, steps=steps,
                           sway_sampling_coef=sway_sampling_coef, seed=seed, max_duration=max_duration, vocoder=vocoder,
                           no_ref_audio=no_ref_audio, duplicate_test=duplicate_test, t_inter=t_inter, edit_mask=edit_mask)

    @torch.no_grad()
    def sample_tts(self, cond, text, duration=None, *, lens=None, steps=32, alpha_spk=1.0, alpha_txt=1.0,
                   sway_sampling_coef=None, seed=None, max_duration=4096, vocoder=None, no_ref_audio=False,
                   duplicate_test=False, t_inter=0.1, edit_mask=None):
        """Reference model/cfm.py:93-223: three-way guidance for TTS on a tripled batch, branch order
        speaker+text | text only | null (drop_audio_cond, drop_text, drop_ppg per cfm.py:166-185); ppg is None."""
        branches = [(False, False, True), (True, False, True), (True, True, True)]
        g = _lib.Guidance(3, float(alpha_spk), float(alpha_txt))
        return self._solve(cond, text, None, duration, branches, g, lens=lens, steps=steps,
                           sway_sampling_coef=sway_sampling_coef, seed=seed, max_duration=max_duration, vocoder=vocoder,
                           no_ref_audio=no_ref_audio, duplicate_test=duplicate_test, t_inter=t_inter, edit_mask=edit_mask)

    @torch.no_grad()
    def sample_vc(self, cond, ppg=None, duration=None, *, lens=None, steps=32, alpha_spk=1.0, alpha_ppg=1.0,
                  sway_sampling_coef=None, seed=None, max_duration=4096, vocoder=None, no_ref_audio=False,
                  duplicate_test=False, t_inter=0.1, edit_mask=None):
        """Reference model/cfm.py:225-346: three-way guidance for voice conversion, branch order
        speaker+ppg | ppg only | null (cfm.py:298-317); text is None (TextEmbedding's zero-token path, dit.py:55-56)."""
        if self.transformer.text_mask_padding and self.transformer.conv_layers > 0:
            # the reference raises NameError here (dit.py:81: text_mask is undefined when text is None)
            raise ValueError("f5e_tts_b200.CFM.sample_vc needs a DiT built with text_mask_padding=False (as the reference)")
        branches = [(False, True, False), (True, True, False), (True, True, True)]
        g = _lib.Guidance(3, float(alpha_spk), float(alpha_ppg))
        return self._solve(cond, None, ppg, duration, branches, g, lens=lens, steps=steps,
                           sway_sampling_coef=sway_sampling_coef, seed=seed, max_duration=max_duration, vocoder=vocoder,
                           no_ref_audio=no_ref_audio, duplicate_test=duplicate_test, t_inter=t_inter, edit_mask=edit_mask)

    def _solve(self, cond, text, ppg, duration, branches, guidance, **kw):
        # The reference's long-form caller runs `model.sample` from a ThreadPoolExecutor (utils_infer.py:503-515); the
        # DiT workspace (plan, caches, CUDA graph) is one per model, so concurrent callers take turns here.
        with _SOLVE_LOCK:
            return self._solve_locked(cond, text, ppg, duration, branches, guidance, **kw)

    def _solve_locked(self, cond, text, ppg, duration, branches, guidance, *, lens, steps, sway_sampling_coef, seed,
                      max_duration, vocoder, no_ref_audio, duplicate_test, t_inter, edit_mask):
        """The part the three samplers of the reference share line for line (cfm.py:119-164 + 187-223, 250-296 + 319-346,
        374-428 + 449-482).  `branches` = (drop_audio_cond, drop_text, drop_ppg) of every stacked evaluation."""
        self.eval()
        dit = self.transformer
        if self.odeint_kwargs.get("method", "euler") != "euler":
            raise NotImplementedError("f5e_tts_b200.CFM: only the fixed-grid Euler solver is implemented")
        dtype = next(self.parameters()).dtype
        if cond.ndim == 2:  # raw wave
            cond = self.mel_spec(cond).permute(0, 2, 1)
            assert cond.shape[-1] == self.num_channels
        cond = cond.to(dtype)
        batch, cond_seq_len, device = *cond.shape[:2], cond.device
        if lens is None:
            lens = torch.full((batch,), cond_seq_len, device=device, dtype=torch.long)
        if isinstance(text, list):
            if self.vocab_char_map is not None:
                if self.use_align_loss or self.use_cross_mask:
                    text = [[x for pair in zip(["_"] * len(t), t) for x in pair] + ["_"] for t in text]  # durpred.intersperse
                text = list_str_to_idx(text, self.vocab_char_map).to(device)
            else:
                text = list_str_to_tensor(text).to(device)
            assert text.shape[0] == batch
        # ---- integer / mask logic: identical ops to cfm.py:393-428 ----
        cond_mask = lens_to_mask(lens)
        if edit_mask is not None:
            cond_mask = cond_mask & edit_mask
        if isinstance(duration, int):
            duration = torch.full((batch,), duration, device=device, dtype=torch.long)
        if text is not None:
            duration = torch.maximum(torch.maximum((text != -1).sum(dim=-1), lens) + 1, duration)
        else:
            duration = torch.maximum(lens + 1, duration)
        duration = duration.clamp(max=max_duration)
        max_duration = int(duration.amax())
        if duplicate_test:
            test_cond = F.pad(cond, (0, 0, cond_seq_len, max_duration - 2 * cond_seq_len), value=0.0)
        cond = F.pad(cond, (0, 0, 0, max_duration - cond_seq_len), value=0.0)
        if no_ref_audio:
            cond = torch.zeros_like(cond)
        cond_mask = F.pad(cond_mask, (0, max_duration - cond_mask.shape[-1]), value=False).unsqueeze(-1)
        step_cond = torch.where(cond_mask, cond, torch.zeros_like(cond))
        # ---- noise: torch's RNG so that y0 is bit-identical to the reference (cfm.py:449-457) ----
        y0 = []
        for dur in duration:
            if seed is not None:
                torch.manual_seed(seed)
            y0.append(torch.randn(int(dur), self.num_channels, device=self.device, dtype=step_cond.dtype))
        y0 = torch.nn.utils.rnn.pad_sequence(y0, padding_value=0, batch_first=True)
        t_start = 0
        if duplicate_test:
            t_start = t_inter
            y0 = (1 - t_start) * y0 + t_start * test_cond
            steps = int(steps * (1 - t_start))
        t = torch.linspace(t_start, 1, steps + 1, device=self.device, dtype=step_cond.dtype)
        if sway_sampling_coef is not None:
            t = t + sway_sampling_coef * (torch.cos(torch.pi / 2 * t) - 1 + t)
        dts = (t[1:] - t[:-1]).float().tolist()  # torchdiffeq: dt = t1 - t0 in the model dtype

        # ---- device loop ----
        P = dit._get_plan()
        B, N, md, td = batch, max_duration, self.num_channels, dit.text_dim
        nb = len(branches)
        S = nb * B
        tens, buf = dit._workspace(S, N)
        xc = tens["x_cat"]
        rows = B * N
        t_emb, p_emb = {}, {}  # the reference caches text_cond / text_uncond the same way (dit.py:433-441)
        isolate = bool(self.isolate_samples) and batch > 1
        P.model.isolate_samples = int(isolate)
        # Folding the LayerNorms into the QKV / FF1 GEMMs costs one rewrite of those weights per Euler step (~0.9 GB of HBM
        # traffic for F5TTS_v1_Base) and saves writing + reading the normalised activations (8 KB per row and layer): it pays
        # from a few thousand rows on, not at the batch-1 latency shape.
        fold = bool(getattr(dit, "fold_layernorm", False)) and S * N >= dit.fold_min_rows
        P.model.fold_jobs = P.fold_jobs.data_ptr() if fold else None
        P.model.fold_n_jobs, P.model.fold_rows = (P.fold_n_jobs, P.fold_rows) if fold else (0, 0)

        def embed(fn, src, drop):  # isolate: every sample embedded alone at its own length, zeros beyond it
            if not isolate:
                return fn(src, N, drop)
            out = torch.zeros(B, N, td, device=xc.device, dtype=torch.bfloat16)
            for i, d in enumerate(duration.tolist()):
                out[i, :d] = fn(None if src is None else src[i:i + 1], d, drop)[0]
            return out
        for i, (drop_audio, drop_text, drop_ppg) in enumerate(branches):
            xh = xc[i * rows:(i + 1) * rows]
            ops.pack_cols(xh, 0, y0.reshape(rows, md))
            ops.pack_cols(xh, md, None if drop_audio else step_cond.reshape(rows, md), md)
            if drop_text not in t_emb:
                t_emb[drop_text] = embed(lambda tx, n, dr: dit._text_embed(tx, B if not isolate else 1, n, dr), text, drop_text)
            ops.pack_cols(xh, 2 * md, t_emb[drop_text].view(rows, td))
            if dit.use_ppg:
                pk = drop_ppg or ppg is None
                if pk not in p_emb:
                    p_emb[pk] = embed(lambda pg, n, dr: dit._ppg_embed(pg, n, dr, B if not isolate else 1), ppg, drop_ppg)
                ops.pack_cols(xh, 2 * md + td, p_emb[pk].view(rows, td))
        del t_emb, p_emb
        mod = dit._modulation(t[:-1])
        klen = duration.to(torch.int32).repeat(nb).contiguous() if batch > 1 else None
        # batch > 1: the DiT blocks run on the valid frames only (no FLOPs on padding)
        want_pack = self.pack_valid_rows
        keep_padded = False
        if want_pack is None:
            want_pack = True
            keep_padded = not (isolate or bool((duration == max_duration).all()))
        pack, pack_keep = (make_packing(klen, N, keep_padded) if (klen is not None and want_pack) else (None, None))
        drops_rows = pack is not None and not keep_padded
        y = y0.float().contiguous()
        traj = None
        if self.store_trajectory:  # a dropping packing writes the valid rows only: padded rows keep y0 there (zeros)
            traj = (torch.zeros if drops_rows else torch.empty)(steps, B, N, md, device=y.device, dtype=torch.float32)
        if self.use_cuda_graph:
            # the cached graph bakes every address in: keep the solve's buffers per shape and copy this call's values in
            gk = (S, N, steps, traj is not None, None if pack is None else (pack.rows_packed, pack.rows_total), klen is None)
            gb = dit._gbuf.get(gk)
            if gb is None:
                if len(dit._gbuf) >= 4:
                    _lib.load().f5e_graph_cache_clear()
                    dit._gbuf.clear()
                gb = dit._gbuf[gk] = dict(y=torch.empty_like(y), mod=torch.empty_like(mod),
                                          traj=None if traj is None else torch.empty_like(traj),
                                          klen=None if klen is None else torch.empty_like(klen),
                                          pack=None if pack is None else [torch.empty_like(t) for t in pack_keep])
            gb["y"].copy_(y)
            gb["mod"].copy_(mod)
            y, mod = gb["y"], gb["mod"]
            if traj is not None:
                if drops_rows:
                    gb["traj"].zero_()
                traj = gb["traj"]
            if klen is not None:
                gb["klen"].copy_(klen)
                klen = gb["klen"]
            if pack is not None:
                for dst, src in zip(gb["pack"], pack_keep):
                    dst.copy_(src)
                pack_keep = gb["pack"]
                pack = _Packing(pack_keep[0].data_ptr(), pack_keep[1].data_ptr(), pack_keep[2].data_ptr(), pack.rows_packed,
                                    pack.rows_total)
        dt_arr = (C.c_float * steps)(*dts)
        check(_lib.load().f5e_ode_solve_guided(
            C.byref(P.model), C.byref(buf), ptr(y), ptr(mod), P.mod_size, dt_arr, steps, C.byref(guidance), ptr(klen), B, N,
            ptr(traj), int(self.emulate_lowp_state and dtype != torch.float32), int(self.use_cuda_graph),
            None if pack is None else C.byref(pack), stream_ptr()), "f5e_ode_solve_guided")
        del pack_keep
        dit.clear_cache()
        sampled = y.to(dtype) if (y.dtype != dtype or not self.use_cuda_graph) else y.clone()
        if traj is not None:
            trajectory = torch.cat([y0.unsqueeze(0), traj.to(dtype)], dim=0)
        else:
            trajectory = torch.stack([y0, sampled])
        out = torch.where(cond_mask, cond, sampled)
        # integer / bool side products of the last solve (tests compare them bit for bit with the reference's)
        self.last_aux = dict(duration=duration, cond_mask=cond_mask, mask=lens_to_mask(duration) if batch > 1 else None, t=t,
                             packed_rows=pack is not None, dropped_padded_rows=drops_rows)
        if vocoder is not None:
            out = vocoder(out.permute(0, 2, 1))
        return out, trajectory
