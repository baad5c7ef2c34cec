"""Per-kernel timings at the BASELINE config-2 shapes (developer tool; prints one JSON line per kernel).

    python tools/microbench.py [--batch 16] [--npad 1030] > gpurun_out/micro.jsonl
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from f5e_tts_b200 import _lib, ops  # noqa: E402

PEAK_TF = 1655.8
PEAK_GBS = 6546.6


ITERS, WARM = 20, 3


def timeit(fn, iters=None, warm=None):
    iters, warm = ITERS if iters is None else iters, WARM if warm is None else warm
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--npad", type=int, default=1030)
    ap.add_argument("--only", default="")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warm", type=int, default=3)
    a = ap.parse_args()
    global ITERS, WARM
    ITERS, WARM = a.iters, a.warm
    dev = torch.device("cuda:0")
    _lib.require_device()
    S, N = 2 * a.batch, a.npad
    M = S * N
    D = 1024
    out = []

    def rec(name, ms, flops=None, bytes_=None, **kw):
        r = dict(kernel=name, ms=round(ms, 4), **kw)
        if flops:
            r["tflops"] = round(flops / ms / 1e9, 1)
            r["frac_tensor"] = round(flops / ms / 1e9 / PEAK_TF, 3)
        if bytes_:
            r["gbs"] = round(bytes_ / ms / 1e6, 1)
            r["frac_hbm"] = round(bytes_ / ms / 1e6 / PEAK_GBS, 3)
        print(json.dumps(r), flush=True)
        out.append(r)

    torch.manual_seed(0)
    if not a.only or "gemm" in a.only:
        for (n, k, act, tag) in [(3072, 1024, 0, "qkv"), (1024, 1024, 0, "out"), (2048, 1024, _lib.ACT_GELU_TANH, "ff1"),
                                 (1024, 2048, 0, "ff2"), (1024, 712, 0, "in_proj"), (100, 1024, 0, "proj_out")]:
            x = torch.randn(M, k, device=dev).bfloat16()
            w = (torch.randn(n, k, device=dev) / 32).bfloat16()
            b = torch.randn(n, device=dev)
            f32 = n % 8 != 0
            o = torch.empty(M, n, device=dev, dtype=torch.float32 if f32 else torch.bfloat16)
            res = o if tag in ("out", "ff2") else None  # in place (x += gate * branch), as in the DiT block
            gate = torch.randn(n, device=dev) if res is not None else None
            ms = timeit(lambda: ops.gemm(x, w, bias=b, act=act, out=o, residual=res, gate=gate, n_pad=N))
            rec(f"gemm_{tag}", ms, flops=2.0 * M * n * k, M=M, N=n, K=k)
            if tag == "qkv":
                inv = 1.0 / (10000 ** (torch.arange(0, 64, 2, device=dev).float() / 64))
                invc = inv.contiguous()
                ms = timeit(lambda: ops.gemm(x, w, bias=b, out=o, n_pad=N, rope=invc, rope_heads=16, rope_inner=1024))
                rec("gemm_qkv_rope_compute", ms, flops=2.0 * M * n * k, M=M, N=n, K=k)
            if res is not None:  # the residual READ in the epilogue + plain store, instead of the in-place TMA reduce-add
                r2 = torch.randn(M, n, device=dev).bfloat16()
                ms = timeit(lambda: ops.gemm(x, w, bias=b, act=act, out=o, residual=r2, gate=gate, n_pad=N))
                rec(f"gemm_{tag}_residual_read", ms, flops=2.0 * M * n * k, M=M, N=n, K=k)
            if tag in ("qkv", "ff1"):
                ms = timeit(lambda: ops.gemm(x, w, out=o, n_pad=N))
                rec(f"gemm_{tag}_plain", ms, flops=2.0 * M * n * k, M=M, N=n, K=k)
            ms = timeit(lambda: torch.matmul(x, w.t()))
            rec(f"cublas_{tag}", ms, flops=2.0 * M * n * k, M=M, N=n, K=k)
    if not a.only or "attn" in a.only:
        qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
        o = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
        klen = torch.full((S,), N, device=dev, dtype=torch.int32)
        ms = timeit(lambda: ops.attention(qkv, 16, klen=klen, n_pad=N, out=o))
        rec("attention", ms, flops=4.0 * S * 16 * N * N * 64, samples=S, n_pad=N)
        q, k, v = qkv.view(S, N, 3, 16, 64).permute(2, 0, 3, 1, 4)
        ms = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v))
        rec("sdpa_torch", ms, flops=4.0 * S * 16 * N * N * 64)
    if not a.only or "conv" in a.only:
        x = torch.randn(M, D, device=dev).bfloat16()
        w = ops.pack_conv_pos_weight((torch.randn(D, 64, 31, device=dev) / 44).bfloat16())
        b = torch.randn(D, device=dev)
        o = torch.empty_like(x)
        ms = timeit(lambda: ops.conv_pos(x, w, b, S, N, out=o))
        rec("conv_pos", ms, flops=2.0 * M * D * 64 * 31)
    if not a.only or "norm" in a.only:
        x = torch.randn(M, D, device=dev).bfloat16()
        mul, add = torch.randn(D, device=dev), torch.randn(D, device=dev)
        o = torch.empty_like(x)
        ms = timeit(lambda: ops.norm_mod(x, mul, add, out=o))
        rec("norm_mod", ms, bytes_=2.0 * M * D * 2)
        import ctypes as C
        lib = _lib.load()
        rstd = torch.empty(M, device=dev)
        a2 = torch.zeros(M, 64, device=dev, dtype=torch.bfloat16)
        st = _lib.stream_ptr()
        ms = timeit(lambda: lib.f5e_row_stats(_lib.ptr(x), _lib.ptr(rstd), _lib.ptr(a2), M, D, C.c_float(1e-6), st))
        rec("row_stats", ms, bytes_=1.0 * M * (D * 2 + 4 + 8))
        pred = torch.randn(2, M // 2, 100, device=dev)
        y = torch.randn(M // 2, 100, device=dev)
        xc = torch.zeros(M, 712, device=dev, dtype=torch.bfloat16)
        ms = timeit(lambda: ops.cfg_euler(pred, y, 0.01, 2.0, x_cat=xc))
        rec("cfg_euler", ms, bytes_=4.0 * (M // 2) * 100 * 4)


if __name__ == "__main__":
    main()
