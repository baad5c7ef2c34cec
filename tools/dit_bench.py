"""Developer tool: CFM.sample at BASELINE config-2 shape with a per-kernel-class time breakdown from the library's
in-stream event profiler.   python tools/dit_bench.py [--batch 16] [--nfe 32] [--solves 2] [--depth 22]"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from f5e_tts_b200 import _lib  # noqa: E402

CLS = {1: "gemm", 2: "attention", 4: "conv_pos", 8: "norm_mod", 16: "elementwise", 32: "dsp", 64: "vq"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--nfe", type=int, default=32)
    ap.add_argument("--solves", type=int, default=2)
    ap.add_argument("--seconds", type=float, default=10.0)
    ap.add_argument("--graph", type=int, default=0)
    ap.add_argument("--prof", type=int, default=1)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    model, _ = bench.build_model(dev)
    model.store_trajectory = False
    model.use_cuda_graph = bool(a.graph)
    w = bench.make_workload(a.batch, a.seconds, 0)
    cond = torch.randn(a.batch, int(w["ref_len"].max()), 100, device=dev)
    text, ref, dur = w["text"].to(dev), w["ref_len"].to(dev), w["dur"].to(dev)
    kw = dict(steps=a.nfe, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)
    model.sample(cond, text, None, dur, lens=ref, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.solves):
        model.sample(cond, text, None, dur, lens=ref, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.solves
    flops = 2.0 * sum(bench.dit_flops(float(n)) for n in w["dur"]) * a.nfe
    res = dict(ms_per_solve=ms, ms_per_nfe_step=ms / a.nfe, audio_s_per_s=bench.audio_seconds(w) / (ms / 1e3),
               model_tflops=flops / ms / 1e9, graph=a.graph, n_pad=int(w["dur"].max()), batch=a.batch)
    if a.prof:
        lib.f5e_prof_begin(127)
        model.sample(cond, text, None, dur, lens=ref, **kw)
        n = lib.f5e_prof_end()
        agg = {}
        cls, ms_, dims = C.c_int32(), C.c_float(), (C.c_int64 * 3)()
        for i in range(n):
            lib.f5e_prof_read(C.c_int64(i), C.byref(cls), C.byref(ms_), dims)
            key = CLS.get(cls.value, str(cls.value))
            if cls.value == 1:
                key += f"_{dims[1]}x{dims[2]}"
            t = agg.setdefault(key, [0.0, 0])
            t[0] += ms_.value
            t[1] += 1
        lib.f5e_prof_begin(0)
        tot = sum(v[0] for v in agg.values())
        res["profile_ms"] = {k: dict(ms=round(v[0], 3), n=v[1], share=round(v[0] / tot, 4)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])}
        res["profile_total_ms"] = tot
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
