"""Kernel-level parity on the GPU, through the C ABI, against plain torch fp32 math of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (300, 1024, 1024), (1000, 3072, 1024), (777, 100, 1024),
                                   (2056, 2048, 712), (33, 512, 256)])
def test_gemm_bias(cuda_dev, M, N, K):
    from f5e_tts_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device=cuda_dev, generator=g).bfloat16()
    w = (torch.randn(N, K, device=cuda_dev, generator=g) / math.sqrt(K)).bfloat16()
    b = torch.randn(N, device=cuda_dev, generator=g)
    f32 = (N % 8 != 0)
    out = ops.gemm(a, w, bias=b, out_f32=f32)
    ref = a.float() @ w.float().t() + b
    torch.cuda.synchronize()
    assert rel_l2(out, ref) < (1e-5 if f32 else 4e-3), rel_l2(out, ref)
    assert (out.float() - ref).abs().max().item() < (1e-3 if f32 else 0.06)


def test_gemm_full_epilogue(cuda_dev):
    from f5e_tts_b200 import ops, _lib
    torch.manual_seed(1)
    B, n_pad, D = 3, 200, 1024
    M = B * n_pad
    a = torch.randn(M, D, device=cuda_dev).bfloat16()
    w = (torch.randn(D, D, device=cuda_dev) / 32).bfloat16()
    bias = torch.randn(D, device=cuda_dev)
    gate = torch.randn(D, device=cuda_dev)
    res = torch.randn(M, D, device=cuda_dev).bfloat16()
    lens = torch.tensor([200, 150, 7], device=cuda_dev, dtype=torch.int32)
    out = ops.gemm(a, w, bias=bias, act=_lib.ACT_GELU_TANH, gate=gate, residual=res, row_len=lens, n_pad=n_pad)
    v = torch.nn.functional.gelu(a.float() @ w.float().t() + bias, approximate="tanh")
    valid = (torch.arange(n_pad, device=cuda_dev)[None] < lens[:, None]).reshape(M, 1)
    ref = res.float() + gate * torch.where(valid, v, torch.zeros_like(v))
    assert rel_l2(out, ref) < 4e-3


def test_gemm_qkv_rope(cuda_dev):
    from f5e_tts_b200 import ops
    torch.manual_seed(2)
    B, n_pad, D, H = 2, 333, 1024, 16
    M = B * n_pad
    a = torch.randn(M, D, device=cuda_dev).bfloat16()
    w = (torch.randn(3 * D, D, device=cuda_dev) / 32).bfloat16()
    bias = torch.randn(3 * D, device=cuda_dev)
    inv = 1.0 / (10000 ** (torch.arange(0, 64, 2, device=cuda_dev).float() / 64))
    ang = torch.arange(n_pad, device=cuda_dev).float()[:, None] * inv[None]
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    for heads, rope in ((16, inv.contiguous()), (1, inv.contiguous())):
        out = ops.gemm(a, w, bias=bias, rope=rope, rope_heads=heads, rope_inner=D, n_pad=n_pad)
        y = (a.float() @ w.float().t() + bias).view(B, n_pad, 3, H, 32, 2)
        c, s = cos[None, :, None, :], sin[None, :, None, :]
        ref = y.clone()
        for part in (0, 1):
            xa, xb = y[:, :, part, :heads, :, 0], y[:, :, part, :heads, :, 1]
            ref[:, :, part, :heads, :, 0] = xa * c - xb * s
            ref[:, :, part, :heads, :, 1] = xb * c + xa * s
        assert rel_l2(out, ref.reshape(M, 3 * D)) < 4e-3


@pytest.mark.parametrize("D", [512, 768, 1024])
def test_norm_mod(cuda_dev, D):
    from f5e_tts_b200 import ops
    torch.manual_seed(3)
    B, n_pad = 4, 123
    x = (torch.randn(B, n_pad, D, device=cuda_dev) * 3 + 1).bfloat16()
    mul = torch.randn(B, D, device=cuda_dev)
    add = torch.randn(B, D, device=cuda_dev)
    ln = torch.nn.functional.layer_norm(x.float(), (D,), eps=1e-6)
    out = ops.norm_mod(x, mul[0], add[0])
    assert rel_l2(out, ln * (1 + mul[0]) + add[0]) < 3e-3
    out = ops.norm_mod(x, mul, add, mod_stride=D, n_pad=n_pad)
    assert rel_l2(out, ln * (1 + mul[:, None]) + add[:, None]) < 3e-3
    out = ops.norm_mod(x, mul[0], add[0], plus_one=0.0)
    assert rel_l2(out, ln * mul[0] + add[0]) < 3e-3


def test_cfg_euler(cuda_dev):
    from f5e_tts_b200 import ops
    torch.manual_seed(4)
    rows, Cc = 1000, 100
    pred = torch.randn(2, rows, Cc, device=cuda_dev)
    y = torch.randn(rows, Cc, device=cuda_dev)
    y0 = y.clone()
    xcat = torch.zeros(2 * rows, 712, device=cuda_dev, dtype=torch.bfloat16)
    traj = torch.empty(rows, Cc, device=cuda_dev)
    ops.cfg_euler(pred, y, 0.03, 2.0, x_cat=xcat, traj=traj)
    ref = y0 + 0.03 * (pred[0] + (pred[0] - pred[1]) * 2.0)
    assert torch.allclose(y, ref, atol=1e-6, rtol=1e-6)
    assert torch.equal(traj, y)
    assert torch.equal(xcat[:rows, :100], y.bfloat16()) and torch.equal(xcat[rows:, :100], y.bfloat16())
    assert xcat[:, 100:].abs().max().item() == 0


def test_guided_euler_three_way(cuda_dev):
    """sample_tts / sample_vc combine (reference cfm.py:184, :316), same operation order in fp32."""
    from f5e_tts_b200 import ops
    torch.manual_seed(14)
    rows, Cc = 777, 100
    pred = torch.randn(3, rows, Cc, device=cuda_dev)
    y = torch.randn(rows, Cc, device=cuda_dev)
    y0 = y.clone()
    xcat = torch.zeros(3 * rows, 356, device=cuda_dev, dtype=torch.bfloat16)
    ops.guided_euler(pred, y, 0.05, 3, 2.5, 3.0, x_cat=xcat)
    ref = y0 + 0.05 * (2.5 * (pred[0] - pred[1]) + 3.0 * (pred[1] - pred[2]) + pred[2])
    assert torch.allclose(y, ref, atol=2e-6, rtol=1e-6)
    for b in range(3):
        assert torch.equal(xcat[b * rows:(b + 1) * rows, :100], y.bfloat16())
    assert xcat[:, 100:].abs().max().item() == 0
    y1 = y0.clone()
    ops.guided_euler(pred, y1, 0.05, 1, 0.0)
    assert torch.allclose(y1, y0 + 0.05 * pred[0], atol=1e-6, rtol=1e-6)


@pytest.mark.parametrize("samples,n_pad,heads,lens", [(1, 128, 1, None), (2, 300, 4, [300, 129]),
                                                       (3, 1031, 16, [1031, 850, 5]), (1, 940, 16, None)])
def test_attention(cuda_dev, samples, n_pad, heads, lens):
    from f5e_tts_b200 import ops
    torch.manual_seed(5)
    inner = heads * 64
    qkv = torch.randn(samples * n_pad, 3 * inner, device=cuda_dev).bfloat16()
    klen = None if lens is None else torch.tensor(lens, device=cuda_dev, dtype=torch.int32)
    out = ops.attention(qkv, heads, klen=klen, n_pad=n_pad)
    q, k, v = qkv.float().view(samples, n_pad, 3, heads, 64).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) * 0.125
    if lens is not None:
        km = torch.arange(n_pad, device=cuda_dev)[None] < klen[:, None]
        s = s.masked_fill(~km[:, None, None, :], float("-inf"))
    ref = (s.softmax(-1) @ v).permute(0, 2, 1, 3).reshape(samples, n_pad, inner)
    o = out.view(samples, n_pad, inner).float()
    for i in range(samples):
        L = n_pad if lens is None else lens[i]
        assert rel_l2(o[i, :L], ref[i, :L]) < 1e-2, (i, rel_l2(o[i, :L], ref[i, :L]))


@pytest.mark.parametrize("samples,n_pad,dim", [(1, 128, 1024), (3, 333, 1024), (2, 200, 768)])
def test_conv_pos(cuda_dev, samples, n_pad, dim):
    from f5e_tts_b200 import ops
    torch.manual_seed(6)
    cpg = dim // 16
    x = torch.randn(samples, n_pad, dim, device=cuda_dev).bfloat16()
    w = (torch.randn(dim, cpg, 31, device=cuda_dev) / math.sqrt(cpg * 31)).bfloat16()
    b = torch.randn(dim, device=cuda_dev)
    res = torch.randn(samples * n_pad, dim, device=cuda_dev).bfloat16()
    out = ops.conv_pos(x.view(-1, dim), ops.pack_conv_pos_weight(w), b, samples, n_pad, residual=res)
    ref = torch.nn.functional.conv1d(x.float().transpose(1, 2), w.float(), b, padding=15, groups=16).transpose(1, 2)
    ref = torch.nn.functional.mish(ref).reshape(-1, dim) + res.float()
    assert rel_l2(out, ref) < 4e-3, rel_l2(out, ref)
