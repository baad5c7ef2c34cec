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


# ---- the fast paths bench.py actually runs, one at a time (VERDICT r01 "weak" 1) -------------------------------------
@pytest.mark.parametrize("M,N,K,residual", [(30536, 3072, 1024, False), (30536, 2048, 1024, False), (30536, 1024, 2048, True),
                                            (30536, 1024, 1024, True), (19077, 3072, 1024, False)])
def test_gemm_cta_pair_tiles(cuda_dev, M, N, K, residual):
    """Shapes that take gemm_bf16_kernel<256, 5, 2> (cta_group::2, >= 74 tiles of 256 x 256), M not a multiple of 256, and the
    N = 1024 GEMMs with the in-place gate * branch + residual (TMA reduce-add) of the DiT block."""
    from f5e_tts_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device=cuda_dev, generator=g).bfloat16()
    w = (torch.randn(N, K, device=cuda_dev, generator=g) / math.sqrt(K)).bfloat16()
    b = torch.randn(N, device=cuda_dev, generator=g)
    if residual:
        x = torch.randn(M, N, device=cuda_dev, generator=g).bfloat16()
        gate = torch.randn(N, device=cuda_dev, generator=g)
        ref = x.float() + gate * (a.float() @ w.float().t() + b)
        out = ops.gemm(a, w, bias=b, gate=gate, residual=x, out=x, n_pad=M)  # in place, as dit.cu does
    else:
        ref = a.float() @ w.float().t() + b
        out = ops.gemm(a, w, bias=b)
    torch.cuda.synchronize()
    assert rel_l2(out, ref) < 4e-3, rel_l2(out, ref)
    # max-abs in bf16 ulps of the result: one rounding of the store, one more of the addend on the reduce-add path
    err = (out.float() - ref).abs()
    assert (err <= 0.016 * ref.abs().clamp_min(1.0)).all(), (err / ref.abs().clamp_min(1.0)).max().item()
    # edge tiles: last rows (M % 256 != 0) and last columns
    assert rel_l2(out[-70:], ref[-70:]) < 4e-3 and rel_l2(out[:, -64:], ref[:, -64:]) < 4e-3


@pytest.mark.parametrize("M,N,outliers", [(1000, 3072, False), (30536, 2048, False), (4096, 3072, True)])
def test_layernorm_folded_gemm(cuda_dev, M, N, outliers):
    """row_stats + fold_ln_weights + the folded GEMM (K = dim + 64, second A tensor map, rstd / d epilogue) against
    F.layer_norm(x) * (1 + scale) + shift -> F.linear in fp32.  `outliers`: rows whose |mean| is 50 x their std and a few
    channels x100 (trained DiTs have such channels): the x . W' - mean * c cancellation must survive them."""
    from f5e_tts_b200 import ops
    D = 1024
    g = torch.Generator(device="cuda").manual_seed(M + N)
    x = torch.randn(M, D, device=cuda_dev, generator=g)
    if outliers:
        x[::3] += 50.0                      # |mean| = 50 std
        x[:, [7, 300, 901]] *= 100.0        # outlier channels
        x[5] = x[5] * 1e-3 + 12.0           # tiny variance around a large mean
    x = x.bfloat16()
    w = (torch.randn(N, D, device=cuda_dev, generator=g) / 32).bfloat16()
    bias = torch.randn(N, device=cuda_dev, generator=g)
    scale = 0.3 * torch.randn(D, device=cuda_dev, generator=g)
    shift = 0.3 * torch.randn(D, device=cuda_dev, generator=g)
    xf = x.float()
    ln = torch.nn.functional.layer_norm(xf, (D,), eps=1e-6)
    ref = (ln * (1 + scale) + shift) @ w.float().t() + bias
    # row_stats in isolation
    rstd, a2 = ops.row_stats(x)
    mean = xf.mean(-1)
    var = xf.var(-1, unbiased=False)
    assert torch.allclose(rstd, torch.rsqrt(var + 1e-6), rtol=2e-4, atol=0)
    m_rec = -(a2[:, 0].float() + a2[:, 1].float())
    assert torch.equal(a2[:, 0], a2[:, 2]) and (a2[:, 3:] == 0).all()
    assert torch.allclose(m_rec, mean, rtol=1e-4, atol=1e-5)
    # fold_ln_weights in isolation
    wf, c, d = ops.fold_ln_weights(w, bias, scale, shift)
    wp = (w.float() * (1 + scale)).bfloat16()
    assert torch.equal(wf[:, :D], wp)
    assert torch.allclose(c, wp.float().sum(-1), rtol=1e-5, atol=1e-3)
    assert torch.allclose(d, bias + w.float() @ shift, rtol=1e-4, atol=1e-3)
    c_rec = wf[:, D].float() + wf[:, D + 2].float()
    assert torch.equal(wf[:, D], wf[:, D + 1]) and torch.allclose(c_rec, c, rtol=1e-4, atol=1e-3) and (wf[:, D + 3:] == 0).all()
    # the folded GEMM against the materialised path of the product (norm_mod -> GEMM) and against fp32
    out = ops.gemm(x, wf, bias=d, ln_rstd=rstd, ln_a2=a2)
    xn = ops.norm_mod(x, scale, shift)
    out_mat = ops.gemm(xn, w, bias=bias)
    e_fold, e_mat = rel_l2(out, ref), rel_l2(out_mat, ref)
    print(f"folded GEMM M={M} N={N} outliers={outliers}: rel-L2 vs fp32 folded {e_fold:.3e}, materialised {e_mat:.3e}")
    assert e_fold < 4e-3 and e_fold < 1.5 * e_mat + 1e-3
    rows_err = (out.float() - ref).norm(dim=-1) / ref.norm(dim=-1)
    assert rows_err.max().item() < 2e-2, rows_err.max().item()   # no single row (e.g. the tiny-variance one) blows up


def test_text_embedder_kernels(cuda_dev):
    """text_gather, dwconv7, grn (ConvNeXtV2 text embedder, backbones/dit.py:54-87, modules.py:225-269) each against torch."""
    from f5e_tts_b200 import ops
    from f5e_tts_b200.tables import text_pos_table
    g = torch.Generator(device="cuda").manual_seed(11)
    B, nt, N, dim = 3, 37, 90, 512
    table = torch.randn(60, dim, device=cuda_dev, generator=g).bfloat16()
    text = torch.randint(0, 59, (B, nt), device=cuda_dev, generator=g)
    text[1, 20:] = -1
    text[2, 5:] = -1
    pos = text_pos_table(dim).to(cuda_dev)
    out = ops.text_gather(text, table, pos, N, drop=False, mask_padding=True, add_pos=True)
    tt = torch.nn.functional.pad(text + 1, (0, N - nt), value=0)          # dit.py:58-63: +1, filler token 0
    ref = table.float()[tt] + pos[:N][None]
    ref = ref.masked_fill((tt == 0)[..., None], 0.0)                      # dit.py:81-84 with mask_padding
    assert rel_l2(out, ref) < 4e-3
    assert (out[1, 20:] == 0).all() and (out[:, nt:] == 0).all()
    outd = ops.text_gather(text, table, pos, N, drop=True, mask_padding=False, add_pos=True)
    assert rel_l2(outd, (table.float()[torch.zeros_like(tt)] + pos[:N][None])) < 4e-3   # drop_text: every token 0
    # depthwise conv k=7 (padding 3) over frames
    x = torch.randn(B, N, dim, device=cuda_dev, generator=g).bfloat16()
    w = torch.randn(dim, 7, device=cuda_dev, generator=g) / 3
    b = torch.randn(dim, device=cuda_dev, generator=g)
    got = ops.dwconv7(x, w, b)
    ref = torch.nn.functional.conv1d(x.float().transpose(1, 2), w[:, None], b, padding=3, groups=dim).transpose(1, 2)
    assert rel_l2(got, ref) < 4e-3
    # GRN (modules.py:225-238): x * (gamma * Nx + 1) + beta, Gx = ||x||_2 over frames, Nx = Gx / mean_c Gx
    C2 = 1024
    x = torch.randn(B, N, C2, device=cuda_dev, generator=g).bfloat16()
    gamma, beta = torch.randn(C2, device=cuda_dev, generator=g), torch.randn(C2, device=cuda_dev, generator=g)
    got = ops.grn(x, gamma, beta)
    xf = x.float()
    Gx = xf.norm(p=2, dim=1, keepdim=True)
    Nx = Gx / (Gx.mean(dim=-1, keepdim=True) + 1e-6)
    ref = gamma * (xf * Nx) + beta + xf
    assert rel_l2(got, ref) < 4e-3
    assert torch.equal(got, ops.grn(x, gamma, beta))  # atomics-free: run-to-run identical


def test_row_movers(cuda_dev):
    """pack_cols, gather_rows, unfold_frames: pure data movement, bit exact against torch indexing."""
    from f5e_tts_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(12)
    rows = 777
    dst = torch.randn(rows, 712, device=cuda_dev, generator=g).bfloat16()
    keep = dst.clone()
    src32 = torch.randn(rows, 100, device=cuda_dev, generator=g)
    srcbf = torch.randn(rows, 512, device=cuda_dev, generator=g).bfloat16()
    ops.pack_cols(dst, 0, src32)
    ops.pack_cols(dst, 100, None, 100)
    ops.pack_cols(dst, 200, srcbf)
    assert torch.equal(dst[:, :100], src32.bfloat16()) and (dst[:, 100:200] == 0).all() and torch.equal(dst[:, 200:], srcbf)
    strided = torch.randn(rows, 300, device=cuda_dev, generator=g)[:, 17:117]    # row-strided fp32 source
    ops.pack_cols(keep, 612, strided)
    assert torch.equal(keep[:, 612:], strided.bfloat16())
    x = torch.randn(1000, 1024, device=cuda_dev, generator=g).bfloat16()
    idx = torch.randperm(1000, device=cuda_dev, generator=g)[:613].to(torch.int32)
    assert torch.equal(ops.gather_rows(x, idx), x[idx.long()])
    B, N, Cc, taps = 2, 57, 64, 5
    x = torch.randn(B, N, Cc, device=cuda_dev, generator=g).bfloat16()
    u = ops.unfold_frames(x, taps)
    xp = torch.nn.functional.pad(x, (0, 0, taps // 2, taps // 2))
    ref = torch.stack([xp[:, t:t + N] for t in range(taps)], dim=2).reshape(B, N, taps * Cc)   # column tap * C + c
    assert torch.equal(u, ref)


@pytest.mark.parametrize("M,K", [(30536, 1024), (19077, 2048)])
def test_gemm_fused_layernorm_statistics(cuda_dev, M, K):
    """The out-projection / FF2 form x += gate * (a @ w^T + b) on the CTA-pair kernel with the LayerNorm statistics of the NEW
    rows taken in the epilogue (residual tile TMA-staged instead of the reduce-add): the result against fp32 torch, and
    ln_finalize(stats) against row_stats of the stored result (same rounded values -> same statistics to fp32 round-off),
    including rows whose mean is far from zero."""
    from f5e_tts_b200 import ops
    N = 1024
    g = torch.Generator(device="cuda").manual_seed(M + K)
    a = torch.randn(M, K, device=cuda_dev, generator=g).bfloat16()
    w = (torch.randn(N, K, device=cuda_dev, generator=g) / math.sqrt(K)).bfloat16()
    b = torch.randn(N, device=cuda_dev, generator=g)
    gate = torch.randn(N, device=cuda_dev, generator=g)
    x0 = torch.randn(M, N, device=cuda_dev, generator=g)
    x0[::5] += 30.0
    x0 = x0.bfloat16()
    lens = torch.tensor([M - 1000], device=cuda_dev, dtype=torch.int32)   # the last 1000 rows are "padded": branch zeroed
    valid = (torch.arange(M, device=cuda_dev) < M - 1000)[:, None]
    ref = x0.float() + gate * torch.where(valid, a.float() @ w.float().t() + b, torch.zeros(1, device=cuda_dev))
    x = x0.clone()
    stats = torch.empty(M, N // 128, 2, device=cuda_dev)
    out = ops.gemm(a, w, bias=b, gate=gate, residual=x, out=x, row_len=lens, n_pad=M, stats_out=stats)
    assert out.data_ptr() == x.data_ptr()
    assert rel_l2(x, ref) < 4e-3
    err = (x.float() - ref).abs()
    assert (err <= 0.008 * ref.abs().clamp_min(1.0)).all()   # ONE rounding now (the reduce-add path has two)
    rstd, a2 = ops.ln_finalize(stats, N)
    rstd_ref, a2_ref = ops.row_stats(x)
    assert torch.allclose(rstd, rstd_ref, rtol=1e-4, atol=0)
    mean, mean_ref = -(a2[:, 0].float() + a2[:, 1].float()), -(a2_ref[:, 0].float() + a2_ref[:, 1].float())
    assert torch.allclose(mean, mean_ref, rtol=1e-5, atol=1e-5)
    # and the unfused (reduce-add) path gives the same rows to one bf16 ulp
    x2 = x0.clone()
    ops.gemm(a, w, bias=b, gate=gate, residual=x2, out=x2, row_len=lens, n_pad=M)
    assert rel_l2(x2, x) < 3e-3
