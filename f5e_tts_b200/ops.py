"""Thin tensor-level wrappers over the C ABI (test hooks and building blocks of the drop-in modules).

torch is used only for device memory and the current stream; every op runs a hand-written sm_100a kernel.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import Epilogue, check, ptr, stream_ptr


def _cuda_bf16(t: torch.Tensor, name: str) -> None:
    if not (t.is_cuda and t.dtype == torch.bfloat16):
        raise RuntimeError(f"{name} must be a CUDA bfloat16 tensor (got {t.device}, {t.dtype}); there is no CPU path")


def gemm(a, w, *, bias=None, act=_lib.ACT_NONE, gate=None, gate_stride=0, residual=None, row_len=None, n_pad=0,
         rope=None, rope_heads=0, rope_inner=0, out=None, out_f32=False, n=None, split3=0, ln_rstd=None, ln_a2=None,
         stats_out=None):
    """epilogue(a[M,K] @ w[N,K]^T).  `a`/`w` may be row-strided views (last dim contiguous).
    ln_rstd / ln_a2 (from `row_stats`): LayerNorm folded in - `w` is then the folded weight [N, K + 64] of `fold_ln_weights`
    and `bias` its `d` vector."""
    _cuda_bf16(a, "a"), _cuda_bf16(w, "w")
    assert a.dim() == 2 and w.dim() == 2 and a.stride(1) == 1 and w.stride(1) == 1
    M, K = a.shape
    N = w.shape[0] if n is None else n
    if ln_rstd is not None:
        assert ln_a2 is not None and w.shape[1] == K + 64 and ln_rstd.dtype == torch.float32 and ln_a2.dtype == torch.bfloat16
        K += 64
    assert w.shape[1] == K
    if out is None:
        out = torch.empty(M, 3 * split3 if split3 else N, device=a.device, dtype=torch.float32 if out_f32 else torch.bfloat16)
    ep = Epilogue()
    ep.out_split3 = split3
    ep.residual_f32 = int(residual is not None and residual.dtype == torch.float32)
    ep.out, ep.ldo, ep.out_f32 = out.data_ptr(), out.stride(0), int(out.dtype == torch.float32)
    ep.bias = 0 if bias is None else bias.data_ptr()
    ep.act = act
    ep.gate = 0 if gate is None else gate.data_ptr()
    ep.gate_stride = gate_stride
    ep.residual = 0 if residual is None else residual.data_ptr()
    ep.ldr = 0 if residual is None else residual.stride(0)
    ep.row_len = 0 if row_len is None else row_len.data_ptr()
    ep.n_pad = n_pad
    if rope is not None:  # the inv_freq vector [32] (x_transformers RotaryEmbedding.inv_freq, dim_head 64)
        assert rope.dtype == torch.float32 and rope.numel() == 32 and rope.is_contiguous()
        ep.rope_inv_freq = rope.data_ptr()
        ep.rope_inner, ep.rope_heads = rope_inner, rope_heads
    if ln_rstd is not None:
        ep.ln_rstd, ep.ln_a2 = ln_rstd.data_ptr(), ln_a2.data_ptr()
    if stats_out is not None:  # fp32 [M, N / 128, 2]: LayerNorm statistics of the rows this GEMM produces (x += ... in place)
        assert stats_out.dtype == torch.float32 and stats_out.shape == (M, N // 128, 2) and stats_out.is_contiguous()
        ep.stats_out, ep.stats_slots = stats_out.data_ptr(), N // 128
    for t in (bias, gate):
        assert t is None or t.dtype == torch.float32
    assert row_len is None or row_len.dtype == torch.int32
    check(_lib.load().f5e_gemm_bf16(ptr(a), a.stride(0), ptr(w), w.stride(0), M, N, K, C.byref(ep), stream_ptr()),
          "f5e_gemm_bf16")
    return out


def norm_mod(x, mul, add, *, plus_one=1.0, eps=1e-6, mod_stride=0, n_pad=0, out=None):
    _cuda_bf16(x, "x")
    assert x.is_contiguous() and mul.dtype == torch.float32 and add.dtype == torch.float32
    D = x.shape[-1]
    M = x.numel() // D
    if out is None:
        out = torch.empty_like(x)
    check(_lib.load().f5e_norm_mod(ptr(x), ptr(out), M, D, ptr(mul), ptr(add), mod_stride, n_pad,
                                   C.c_float(plus_one), C.c_float(eps), stream_ptr()), "f5e_norm_mod")
    return out


def cfg_euler(pred, y, dt, cfg, *, x_cat=None, traj=None, round_bf16=False):
    """In-place y += dt * (pc + (pc - pu) * cfg); pred is fp32 [2, rows, ld]."""
    assert pred.dtype == torch.float32 and y.dtype == torch.float32 and y.is_contiguous()
    rows, Cc = y.numel() // y.shape[-1], y.shape[-1]
    ld_pred = pred.stride(-2)
    check(_lib.load().f5e_cfg_euler(ptr(pred), ld_pred, ptr(y), ptr(x_cat), 0 if x_cat is None else x_cat.stride(-2),
                                    ptr(traj), rows, Cc, C.c_float(dt), C.c_float(cfg), int(round_bf16),
                                    stream_ptr()), "f5e_cfg_euler")
    return y


def guided_euler(pred, y, dt, branches, c0, c1=0.0, *, x_cat=None, traj=None, round_bf16=False):
    """In-place Euler step under 1/2/3-branch guidance (f5e_guided_euler); pred is fp32 [branches, rows, ld]."""
    assert pred.dtype == torch.float32 and y.dtype == torch.float32 and y.is_contiguous()
    rows, Cc = y.numel() // y.shape[-1], y.shape[-1]
    g = _lib.Guidance(int(branches), float(c0), float(c1))
    check(_lib.load().f5e_guided_euler(ptr(pred), pred.stride(-2), ptr(y), ptr(x_cat),
                                       0 if x_cat is None else x_cat.stride(-2), ptr(traj), rows, Cc, C.c_float(dt),
                                       C.byref(g), int(round_bf16), stream_ptr()), "f5e_guided_euler")
    return y


def attention(qkv, heads, *, klen=None, n_pad=None, scale=None, out=None):
    """qkv: bf16 [samples*n_pad, 3*heads*64] -> bf16 [samples*n_pad, heads*64]."""
    _cuda_bf16(qkv, "qkv")
    rows = qkv.shape[0]
    inner = heads * 64
    assert qkv.shape[1] == 3 * inner and qkv.stride(1) == 1
    n_pad = rows if n_pad is None else n_pad
    samples = rows // n_pad
    if out is None:
        out = torch.zeros(rows, inner, device=qkv.device, dtype=torch.bfloat16)
    assert klen is None or (klen.dtype == torch.int32 and klen.numel() == samples)
    check(_lib.load().f5e_attention(ptr(qkv), qkv.stride(0), ptr(out), out.stride(0), ptr(klen), samples, n_pad, heads,
                                    C.c_float(scale if scale is not None else 0.125), stream_ptr()), "f5e_attention")
    return out


def pack_conv_pos_weight(w: torch.Tensor) -> torch.Tensor:
    """[dim, cpg, 31] conv weight -> bf16 [31*dim, 64] (tap-major, input channels zero-padded to 64)."""
    dim, cpg, taps = w.shape
    wp = torch.zeros(taps, dim, 64, device=w.device, dtype=torch.bfloat16)
    wp[:, :, :cpg] = w.permute(2, 0, 1).to(torch.bfloat16)
    return wp.reshape(taps * dim, 64).contiguous()


def conv_pos(x, w_packed, bias, samples, n_pad, *, groups=16, act=_lib.ACT_MISH, residual=None, out=None):
    _cuda_bf16(x, "x")
    assert x.is_contiguous() and x.dim() == 2 and x.shape[0] == samples * n_pad
    dim = x.shape[1]
    if out is None:
        out = torch.empty_like(x)
    ep = Epilogue()
    ep.out, ep.ldo = out.data_ptr(), out.stride(0)
    ep.bias = bias.data_ptr()
    ep.act = act
    ep.residual = 0 if residual is None else residual.data_ptr()
    ep.ldr = 0 if residual is None else residual.stride(0)
    check(_lib.load().f5e_conv_pos(ptr(x), ptr(w_packed), samples, n_pad, dim, groups, C.byref(ep), stream_ptr()),
          "f5e_conv_pos")
    return out


def pack_cols(dst, col0, src, cols=None):
    """dst[:, col0:col0+cols] = bf16(src) (src fp32/bf16 2-D, row-strided ok; None writes zeros)."""
    _cuda_bf16(dst, "dst")
    assert dst.dim() == 2 and dst.stride(1) == 1
    rows = dst.shape[0]
    if src is not None:
        if src.dtype not in (torch.float32, torch.bfloat16):
            src = src.float()
        assert src.dim() == 2 and src.shape[0] == rows and src.stride(1) == 1
        cols = src.shape[1] if cols is None else cols
    check(_lib.load().f5e_pack_cols(ptr(dst), dst.stride(0), col0, ptr(src), int(src is not None and src.dtype == torch.float32),
                                    0 if src is None else src.stride(0), rows, cols, stream_ptr()), "f5e_pack_cols")
    return dst


def text_gather(text, table, pos, n_pad, *, drop=False, mask_padding=True, add_pos=True):
    """text int64 [B, nt] (pad -1) -> bf16 [B, n_pad, dim]."""
    _cuda_bf16(table, "table")
    B, nt = text.shape
    text = text.to(torch.int64).contiguous()
    dim = table.shape[1]
    out = torch.empty(B, n_pad, dim, device=table.device, dtype=torch.bfloat16)
    check(_lib.load().f5e_text_gather(ptr(text), B, nt, ptr(table), ptr(pos), 0 if pos is None else pos.shape[0], ptr(out),
                                      n_pad, dim, int(drop), int(mask_padding), int(add_pos and pos is not None),
                                      stream_ptr()), "f5e_text_gather")
    return out


def dwconv7(x, w, b):
    _cuda_bf16(x, "x")
    B, N, Cc = x.shape
    out = torch.empty_like(x)
    check(_lib.load().f5e_dwconv7(ptr(x), ptr(w), ptr(b), ptr(out), B, N, Cc, stream_ptr()), "f5e_dwconv7")
    return out


def grn(x, gamma, beta):
    _cuda_bf16(x, "x")
    B, N, Cc = x.shape
    out = torch.empty_like(x)
    scratch = torch.empty(B, (N + 63) // 64, Cc, device=x.device, dtype=torch.float32)  # slab partials
    check(_lib.load().f5e_grn(ptr(x), ptr(gamma), ptr(beta), ptr(out), ptr(scratch), B, N, Cc, stream_ptr()), "f5e_grn")
    return out


def unfold_frames(x, taps):
    _cuda_bf16(x, "x")
    B, N, Cc = x.shape
    out = torch.empty(B, N, taps * Cc, device=x.device, dtype=torch.bfloat16)
    check(_lib.load().f5e_unfold_frames(ptr(x), ptr(out), B, N, Cc, taps, stream_ptr()), "f5e_unfold_frames")
    return out


def row_stats(x, eps=1e-6):
    """LayerNorm statistics of bf16 x [M, D] for the folded path -> (rstd fp32 [M], a2 bf16 [M, 64] = -mean as hi | lo | hi)."""
    _cuda_bf16(x, "x")
    assert x.dim() == 2 and x.is_contiguous()
    M, D = x.shape
    rstd = torch.empty(M, device=x.device, dtype=torch.float32)
    a2 = torch.zeros(M, 64, device=x.device, dtype=torch.bfloat16)
    check(_lib.load().f5e_row_stats(ptr(x), ptr(rstd), ptr(a2), M, D, C.c_float(eps), stream_ptr()), "f5e_row_stats")
    return rstd, a2


def fold_ln_weights(w, bias, scale, shift):
    """One job of f5e_fold_ln_weights: w bf16 [N, D], bias fp32 [N] or None, scale / shift fp32 [D] ->
    (wf bf16 [N, D + 64], c fp32 [N], d fp32 [N]) with (LN(x) (1 + scale) + shift) @ w^T + bias == rstd (x . wf - mean c) + d."""
    from .model import _FoldJob

    _cuda_bf16(w, "w")
    N, D = w.shape
    dev = w.device
    wf = torch.zeros(N, D + 64, device=dev, dtype=torch.bfloat16)
    c, d = torch.empty(N, device=dev), torch.empty(N, device=dev)
    mod = torch.cat([scale.float(), shift.float()]).contiguous()
    job = _FoldJob(w.data_ptr(), 0 if bias is None else bias.data_ptr(), wf.data_ptr(), c.data_ptr(), d.data_ptr(), N, 0, 0, D)
    jobs = torch.frombuffer(bytearray(bytes(job)), dtype=torch.uint8).to(dev)
    check(_lib.load().f5e_fold_ln_weights(ptr(jobs), 1, N, D, ptr(mod), stream_ptr()), "f5e_fold_ln_weights")
    return wf, c, d


def gather_rows(src, row_map):
    _cuda_bf16(src, "src")
    assert src.dim() == 2 and src.is_contiguous() and row_map.dtype == torch.int32
    out = torch.empty(row_map.numel(), src.shape[1], device=src.device, dtype=torch.bfloat16)
    check(_lib.load().f5e_gather_rows(ptr(src), ptr(out), ptr(row_map), row_map.numel(), src.shape[1], stream_ptr()),
          "f5e_gather_rows")
    return out


def ln_finalize(stats, D, eps=1e-6):
    """(mean, M2) partials [M, D / 128, 2] of gemm(..., stats_out=) -> (rstd [M], a2 [M, 64]) as `row_stats` returns them."""
    M, slots, _ = stats.shape
    rstd = torch.empty(M, device=stats.device, dtype=torch.float32)
    a2 = torch.zeros(M, 64, device=stats.device, dtype=torch.bfloat16)
    check(_lib.load().f5e_ln_finalize(ptr(stats), slots, M, D, C.c_float(eps), ptr(rstd), ptr(a2), stream_ptr()), "f5e_ln_finalize")
    return rstd, a2
