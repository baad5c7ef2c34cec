"""Micro-benchmark of the tcgen05 GEMM at the DiT shapes (CUDA events, L2 flushed between iterations)."""
import sys, os, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from f5e_tts_b200 import ops, _lib

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 30080
for (N, K, name) in [(3072, 1024, "qkv"), (1024, 1024, "out"), (2048, 1024, "ff1"), (1024, 2048, "ff2"), (100, 1024, "proj_out"), (1024, 712, "in")]:
    a = torch.randn(M, K, device=dev).bfloat16()
    w = (torch.randn(N, K, device=dev) / math.sqrt(K)).bfloat16()
    b = torch.randn(N, device=dev)
    f32 = N % 8 != 0
    out = torch.empty(M, N, device=dev, dtype=torch.float32 if f32 else torch.bfloat16)
    for _ in range(3):
        ops.gemm(a, w, bias=b, out=out)
    ts = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.gemm(a, w, bias=b, out=out); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): torch.matmul(a, w.t())
    e1.record(); torch.cuda.synchronize()
    tc = e0.elapsed_time(e1) / 10
    print(f"{name:9s} M={M} N={N} K={K}: {t*1e3:8.1f} us  {2*M*N*K/t/1e9:7.1f} TFLOP/s   (cuBLAS back-to-back {tc*1e3:8.1f} us {2*M*N*K/tc/1e9:7.1f})")
