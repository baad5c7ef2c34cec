"""Developer tool: board power, SM clock and energy per launch of each hot kernel run back to back for ~2 s
(the DiT step is power-capped on a B200: what a step costs is joules, not cycles).

    python tools/power_probe.py > gpurun_out/power.jsonl

For every kernel: launches in a loop for --seconds, nvidia-smi power.draw / clocks.sm sampled every 50 ms (first 0.5 s
dropped), time per launch from CUDA events -> energy per launch = mean power x time per launch.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from f5e_tts_b200 import _lib, ops  # noqa: E402


class Sampler:
    def __init__(self):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=power.draw,clocks.sm,clocks_event_reasons.sw_power_cap",
                                   "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f, stderr=subprocess.DEVNULL)
        self.t0 = time.time()

    def stop(self, drop=10):
        self.p.terminate()
        self.p.wait(timeout=5)
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        rows = rows[drop:]
        pw = [float(r[0]) for r in rows]
        ck = [float(r[1]) for r in rows]
        cap = sum(1 for r in rows if r[2].lower().startswith("active"))
        return statistics.mean(pw), statistics.median(ck), cap / max(1, len(rows)), len(rows)


def probe(name, fn, seconds, work=None):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    # calibrate launches for ~seconds
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        fn()
    b.record()
    torch.cuda.synchronize()
    per = a.elapsed_time(b) / 20
    n = max(20, int(seconds * 1e3 / per))
    s = Sampler()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n
    pw, ck, cap, ns = s.stop()
    r = dict(kernel=name, ms=round(ms, 4), watts=round(pw, 1), sm_mhz=ck, power_capped=round(cap, 2), samples=ns,
             mj_per_launch=round(pw * ms, 2))
    if work:
        r["pj_per_flop"] = round(pw * ms * 1e-3 / work * 1e12, 3)
    print(json.dumps(r), flush=True)
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=2.0)
    ap.add_argument("--only-gemm", action="store_true", help="the four block GEMMs only (+ zero-operand variants)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    _lib.require_device()
    S, N, D = 32, 1030, 1024
    M = S * N
    torch.manual_seed(0)
    # idle (clocked, no work): sleep
    s = Sampler()
    time.sleep(a.seconds)
    pw, ck, cap, ns = s.stop()
    print(json.dumps(dict(kernel="idle", watts=round(pw, 1), sm_mhz=ck)), flush=True)
    for (n, k, act, tag) in [(3072, 1024, 0, "qkv"), (1024, 1024, 0, "out"), (2048, 1024, _lib.ACT_GELU_TANH, "ff1"), (1024, 2048, 0, "ff2")]:
        x = torch.randn(M, k, device=dev).bfloat16()
        w = (torch.randn(n, k, device=dev) / 32).bfloat16()
        b = torch.randn(n, device=dev)
        o = torch.empty(M, n, device=dev, dtype=torch.bfloat16)
        res = o if tag in ("out", "ff2") else None
        gate = torch.randn(n, device=dev) if res is not None else None
        probe(f"gemm_{tag}", lambda: ops.gemm(x, w, bias=b, act=act, out=o, residual=res, gate=gate, n_pad=N), a.seconds, 2.0 * M * n * k)
        if tag == "qkv" and not a.only_gemm:
            probe("cublas_qkv", lambda: torch.matmul(x, w.t()), a.seconds, 2.0 * M * n * k)
        if a.only_gemm and tag in ("qkv", "ff2"):  # how much of the energy is data-dependent switching: all-zero activations
            x0 = torch.zeros_like(x)
            probe(f"gemm_{tag}_zeroA", lambda: ops.gemm(x0, w, bias=b, act=act, out=o, residual=res, gate=gate, n_pad=N), a.seconds, 2.0 * M * n * k)
    if a.only_gemm:
        return
    qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
    o = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
    klen = torch.full((S,), N, device=dev, dtype=torch.int32)
    probe("attention", lambda: ops.attention(qkv, 16, klen=klen, n_pad=N, out=o), a.seconds, 4.0 * S * 16 * N * N * 64)
    q, k, v = qkv.view(S, N, 3, 16, 64).permute(2, 0, 3, 1, 4)
    probe("sdpa_torch", lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v), a.seconds, 4.0 * S * 16 * N * N * 64)
    x = torch.randn(M, D, device=dev).bfloat16()
    w = ops.pack_conv_pos_weight((torch.randn(D, 64, 31, device=dev) / 44).bfloat16())
    b = torch.randn(D, device=dev)
    o2 = torch.empty_like(x)
    probe("conv_pos", lambda: ops.conv_pos(x, w, b, S, N, out=o2), a.seconds, 2.0 * M * D * 64 * 31)
    mul, add = torch.randn(D, device=dev), torch.randn(D, device=dev)
    probe("norm_mod", lambda: ops.norm_mod(x, mul, add, out=o2), a.seconds)
    big = torch.empty(1 << 28, device=dev, dtype=torch.bfloat16)
    big2 = torch.empty_like(big)
    probe("copy_512MB", lambda: big2.copy_(big), a.seconds)


if __name__ == "__main__":
    main()
