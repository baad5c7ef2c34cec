"""GumbelVectorQuantizer mirror (reference model/modules.py:744-950), eval-mode forward on libf5e_b200 (f5e_vq_quantize).

Same constructor, parameter names (`vars`, `weight_proj.{weight,bias}`) and result dict as the reference.  Training-mode
gumbel sampling stays with the reference: this path is the inference-time quantize call (`forward(x, produce_targets=True)`
/ `forward_idx(x)` in eval mode), where the reference's one-hot * vars product reduces to argmax + gather.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from . import _lib
from ._lib import check, ptr, stream_ptr
from .vocos import split3_weight


class GumbelVectorQuantizer(nn.Module):
    def __init__(self, dim, num_vars, temp, groups, combine_groups, vq_dim, time_first, activation=nn.GELU(),
                 weight_proj_depth=1, weight_proj_factor=1, hard=True, std=0):
        super().__init__()
        if weight_proj_depth != 1:
            raise NotImplementedError("f5e_tts_b200.GumbelVectorQuantizer: weight_proj_depth > 1 is not on the F5E path "
                                      "(configs/example.yaml:66 uses 1)")
        self.groups, self.combine_groups, self.input_dim, self.num_vars = groups, combine_groups, dim, num_vars
        self.time_first, self.hard = time_first, hard
        assert vq_dim % groups == 0, f"dim {vq_dim} must be divisible by groups {groups} for concatenation"
        var_dim = vq_dim // groups
        num_groups = groups if not combine_groups else 1
        self.vars = nn.Parameter(torch.empty(1, num_groups * num_vars, var_dim))
        if std == 0:
            nn.init.uniform_(self.vars)
        else:
            nn.init.normal_(self.vars, mean=0, std=std)
        self.weight_proj = nn.Linear(dim, groups * num_vars)
        nn.init.normal_(self.weight_proj.weight, mean=0, std=1)
        nn.init.zeros_(self.weight_proj.bias)
        if isinstance(temp, str):
            import ast

            temp = ast.literal_eval(temp)
        assert len(temp) == 3, f"{temp}, {len(temp)}"
        self.max_temp, self.min_temp, self.temp_decay = temp
        self.curr_temp = self.max_temp
        self.codebook_indices = None
        self.refine_tol = 1e-4      # top-2 margin below which an index is re-decided in sequential fp32
        self.fused_decision = True  # arg-max / gather in the logit GEMM's epilogue (1-2 groups of <= 128 entries); False: unfused
        self.last_refined = None    # device int32 scalar: number of such re-decisions in the last call
        self._plan = None

    # ---- host-side helpers, as in the reference -------------------------------------------------------
    def set_num_updates(self, num_updates):
        self.curr_temp = max(self.max_temp * self.temp_decay ** num_updates, self.min_temp)

    def get_codebook_indices(self):
        if self.codebook_indices is None:
            from itertools import product

            inds = list(product(*([range(self.num_vars)] * self.groups)))
            ci = torch.tensor(inds, dtype=torch.long, device=self.vars.device).flatten()
            if not self.combine_groups:
                ci = ci.view(self.num_vars ** self.groups, -1)
                for b in range(1, self.groups):
                    ci[:, b] += self.num_vars * b
                ci = ci.flatten()
            self.codebook_indices = ci
        return self.codebook_indices

    def codebook(self):
        return self.vars.squeeze(0).index_select(0, self.get_codebook_indices()).view(self.num_vars ** self.groups, -1)

    def to_codebook_index(self, indices):
        res = indices.new_full(indices.shape[:-1], 0)
        for i in range(self.groups):
            res += indices[..., i] * (self.num_vars ** (self.groups - i - 1))
        return res

    def forward_idx(self, x):
        res = self.forward(x, produce_targets=True)
        return res["x"], res["targets"]

    def _apply(self, fn, *a, **k):
        self._plan = None
        return super()._apply(fn, *a, **k)

    def _load_from_state_dict(self, *a, **k):  # also reached through a parent's load_state_dict (DiT.quantizer)
        self._plan = None
        return super()._load_from_state_dict(*a, **k)

    def _param_versions(self):  # in-place parameter updates bump these
        return tuple(p._version for p in self.parameters())

    def _get_plan(self):
        if self._plan is not None and self._plan["versions"] != self._param_versions():
            self._plan = None
        if self._plan is None:
            dev = self.vars.device
            w = self.weight_proj.weight.detach().float().contiguous()
            b = self.weight_proj.bias.detach().float().contiguous()
            G, V = self.groups, self.num_vars
            fused = None
            if G <= 2 and V <= 128:  # fused decision: every group padded to 128 logit columns (f5e_vq_quantize_fused)
                wp = torch.zeros(256, w.shape[1], device=w.device)
                bp = torch.full((256,), -1e30, device=w.device)
                for g in range(G):
                    wp[128 * g: 128 * g + V] = w[g * V: (g + 1) * V]
                    bp[128 * g: 128 * g + V] = b[g * V: (g + 1) * V]
                fused = (split3_weight(wp), bp.contiguous())
            self._plan = dict(w3=split3_weight(w), w=w, b=b, fused=fused,
                              vars=self.vars.detach().float().reshape(-1, self.vars.shape[-1]).contiguous(), dev=dev,
                              versions=self._param_versions())
        return self._plan

    @torch.no_grad()
    def forward(self, x, produce_targets=False):
        if self.training:
            raise NotImplementedError("f5e_tts_b200.GumbelVectorQuantizer is the eval-mode quantize call; call .eval() "
                                      "(gumbel-softmax training stays with the reference)")
        if not x.is_cuda:
            raise RuntimeError("f5e_tts_b200.GumbelVectorQuantizer has no CPU path: pass a CUDA tensor")
        _lib.require_device()
        P = self._get_plan()
        if not self.time_first:
            x = x.transpose(1, 2)
        bsz, tsz, fsz = x.shape
        assert fsz == self.input_dim
        in_dtype = x.dtype
        xf = x.reshape(-1, fsz).float().contiguous()
        M, G, V, vd = bsz * tsz, self.groups, self.num_vars, self.vars.shape[-1]
        dev = x.device
        ldl = (G * V + 3) // 4 * 4
        a3 = torch.empty(M, 3 * fsz, device=dev, dtype=torch.bfloat16)
        targets = torch.empty(bsz, tsz, G, device=dev, dtype=torch.int64)
        out_bf16 = in_dtype == torch.bfloat16
        xq = torch.empty(bsz, tsz, G * vd, device=dev, dtype=torch.bfloat16 if out_bf16 else torch.float32)
        sums = torch.empty(G * V, device=dev, dtype=torch.float32)
        hist = torch.empty(G * V, device=dev, dtype=torch.int32)
        refined = torch.empty(1, device=dev, dtype=torch.int32)
        ppl = torch.empty(2, device=dev, dtype=torch.float32)
        if P["fused"] is not None and self.fused_decision:
            ties = torch.empty(M * G, 2, device=dev, dtype=torch.int32)
            check(_lib.load().f5e_vq_quantize_fused(ptr(xf), fsz, M, fsz, ptr(P["fused"][0]), ptr(P["fused"][1]), ptr(P["w"]),
                                                    ptr(P["b"]), ptr(P["vars"]), G, V, vd, int(self.combine_groups),
                                                    C.c_float(self.refine_tol), ptr(a3), ptr(ties), ptr(targets), ptr(xq),
                                                    int(out_bf16), ptr(sums), ptr(hist), ptr(refined), ptr(ppl), stream_ptr()),
                  "f5e_vq_quantize_fused")
        else:
            logits = torch.empty(M, ldl, device=dev, dtype=torch.float32)
            check(_lib.load().f5e_vq_quantize(ptr(xf), fsz, M, fsz, ptr(P["w3"]), ptr(P["w"]), ptr(P["b"]), ptr(P["vars"]), G, V, vd,
                                              int(self.combine_groups), C.c_float(self.refine_tol), ptr(a3), ptr(logits), ldl,
                                              ptr(targets), ptr(xq), int(out_bf16), ptr(sums), ptr(hist), ptr(refined), ptr(ppl),
                                              stream_ptr()), "f5e_vq_quantize")
        self.last_refined = refined
        if not out_bf16 and in_dtype != torch.float32:
            xq = xq.to(in_dtype)
        result = {"num_vars": V * G, "code_perplexity": ppl[0], "prob_perplexity": ppl[1], "temp": self.curr_temp}
        if produce_targets:
            result["targets"] = targets
        if not self.time_first:
            xq = xq.transpose(1, 2)
        result["x"] = xq
        return result
