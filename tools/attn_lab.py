"""Developer tool: time (and optionally trace) the attention kernel of one library variant at the config-2 shape.

    F5E_LIB_SUFFIX=trace F5E_NVCC_FLAGS=-DATT_TRACE python -m f5e_tts_b200.build     # here (cross-compile)
    F5E_LIB_SUFFIX=trace python tools/attn_lab.py --trace gpurun_out/trace.npy        # on the GPU box

Prints one JSON line: ms per launch (CUDA events, 20 launches after 3 warm-ups), max-abs error vs an fp32 softmax on a
sub-sample, and torch SDPA's time on the same tensors.
"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from f5e_tts_b200 import _lib, ops  # noqa: E402


def timeit(fn, iters=20, warm=3):
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
    ap.add_argument("--samples", type=int, default=32)
    ap.add_argument("--npad", type=int, default=1030)
    ap.add_argument("--ragged", action="store_true", help="per-sample key lengths uniform in [npad/3, npad]")
    ap.add_argument("--trace", default="")
    ap.add_argument("--sdpa", action="store_true")
    ap.add_argument("--sustain", type=float, default=0.0, help="also time the kernel back to back for this many seconds "
                    "(power-capped clocks, what the DiT step sees) with board power sampled from nvidia-smi")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    _lib.require_device()
    S, N, H, D = a.samples, a.npad, 16, 1024
    M = S * N
    torch.manual_seed(0)
    qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
    o = torch.zeros(M, D, device=dev, dtype=torch.bfloat16)
    if a.ragged:
        klen = torch.randint(N // 3, N + 1, (S,), device=dev, dtype=torch.int32)
        klen[0] = N
    else:
        klen = torch.full((S,), N, device=dev, dtype=torch.int32)
    ms = timeit(lambda: ops.attention(qkv, H, klen=klen, n_pad=N, out=o))
    # parity on two samples / three heads vs fp32 softmax
    err = 0.0
    for s_ in (0, S - 1):
        L = int(klen[s_])
        blk = qkv[s_ * N:s_ * N + L].float().view(L, 3, H, 64)
        for h in (0, 7, 15):
            q, k, v = blk[:, 0, h], blk[:, 1, h], blk[:, 2, h]
            ref = torch.softmax(q @ k.t() * 0.125, -1) @ v
            got = o[s_ * N:s_ * N + L].view(L, H, 64)[:, h].float()
            err = max(err, (got - ref).abs().max().item())
    flops = 4.0 * 16 * 64 * float((klen.double() ** 2).sum())
    r = dict(variant=os.environ.get("F5E_LIB_SUFFIX", "product"), ms=round(ms, 4), tflops=round(flops / ms / 1e9, 1),
             max_abs_err=round(err, 5), samples=S, npad=N, ragged=a.ragged)
    if a.sustain > 0:
        import subprocess
        import tempfile
        n = max(20, int(a.sustain * 1e3 / ms))
        f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=power.draw,clocks.sm", "--format=csv,noheader,nounits",
                              "-lms", "50"], stdout=f, stderr=subprocess.DEVNULL)
        sus = timeit(lambda: ops.attention(qkv, H, klen=klen, n_pad=N, out=o), iters=n, warm=1)
        p.terminate()
        p.wait(timeout=5)
        rows = [x.strip().split(", ") for x in open(f.name) if x.strip()][8:]
        os.unlink(f.name)
        r["sustained_ms"] = round(sus, 4)
        if rows:
            r["watts"] = round(sum(float(x[0]) for x in rows) / len(rows), 1)
            r["sm_mhz"] = sorted(float(x[1]) for x in rows)[len(rows) // 2]
            r["mj_per_launch"] = round(r["watts"] * sus, 1)
    if a.sdpa:
        q, k, v = qkv.view(S, N, 3, H, 64).permute(2, 0, 3, 1, 4)
        r["sdpa_ms"] = round(timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v)), 4)
    if a.trace:
        lib = _lib.load()
        nblk, nslot, nrole, ncta = 48, 8, 5, 296  # roles: softmax quarters 0-3, MMA issuer 4
        buf = torch.zeros(ncta * nrole * nblk * nslot, device=dev, dtype=torch.int64)
        assert lib.f5e_attention_trace(C.c_void_p(buf.data_ptr())) == 0
        ops.attention(qkv, H, klen=klen, n_pad=N, out=o)
        torch.cuda.synchronize()
        lib.f5e_attention_trace(C.c_void_p(0))
        import numpy as np
        np.save(a.trace, buf.cpu().numpy().reshape(ncta, nrole, nblk, nslot))
    print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
