"""Developer tool: A/B of programmatic dependent launch (F5E_PDL=0 / 1, read once per process by the library).

    F5E_PDL=0 python tools/pdl_ab.py; F5E_PDL=1 python tools/pdl_ab.py

Prints one JSON line: batch-1 latency shape (stream and cached graph), the config-2 shape (16 x ~10 s, NFE 32, stream and
graph), and an md5 of each result - the two settings must agree bit for bit (same kernels, same arithmetic; only the launch
overlap differs, so any difference is a missing pdl_wait).
"""
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402


def md5(t):
    return hashlib.md5(t.detach().float().cpu().numpy().tobytes()).hexdigest()[:12]


def timed(fn, warm, reps):
    for _ in range(warm):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def main():
    dev = torch.device("cuda:0")
    model, _ = bench.build_model(dev)
    model.store_trajectory = False
    res = dict(pdl=os.environ.get("F5E_PDL", "(default)"))
    g = torch.Generator().manual_seed(7)
    cond = torch.randn(1, 315, 100, generator=g).to(dev)
    text = torch.randint(1, 2545, (1, 150), generator=g).to(dev)
    kw = dict(steps=32, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)
    for graph in (0, 1):
        model.use_cuda_graph = bool(graph)
        ms, (out, _) = timed(lambda: model.sample(cond, text, None, 940, **kw), 2, 4)
        res["b1_%s_ms_per_nfe" % ("graph" if graph else "stream")] = round(ms / 32, 4)
        res["b1_%s_md5" % ("graph" if graph else "stream")] = md5(out)
    if "--sweep" in sys.argv:  # where does it stop paying?  equal-length batches of 940 frames (rows = 2 * B * 940 with CFG)
        for B in (2, 3, 4, 6, 8):
            condb, textb = cond.expand(B, -1, -1).contiguous(), text.expand(B, -1).contiguous()
            for graph in (0, 1):
                model.use_cuda_graph = bool(graph)
                ms, _ = timed(lambda: model.sample(condb, textb, None, 940, **kw), 2, 3)
                res["b%d_%s_ms_per_nfe" % (B, "graph" if graph else "stream")] = round(ms / 32, 4)
        print(json.dumps(res), flush=True)
        return
    w = bench.make_workload(16, 10.0, 0)
    ref_len, dur, text = w["ref_len"].to(dev), w["dur"].to(dev), w["text"].to(dev)
    cond = torch.randn(16, int(ref_len.max()), 100, generator=g).to(dev)
    for graph in (0, 1):
        model.use_cuda_graph = bool(graph)
        ms, (out, _) = timed(lambda: model.sample(cond, text, None, dur, lens=ref_len, **kw), 2, 3)
        res["c2_%s_ms_per_nfe" % ("graph" if graph else "stream")] = round(ms / 32, 4)
        res["c2_%s_md5" % ("graph" if graph else "stream")] = md5(out)
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
