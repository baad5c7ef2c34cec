"""PPG extractor throughput: audio -> kaldi fbank (+CMVN) -> 12-block Conformer -> content embedding.
Workload: configs[2]-shaped batch (32 source utterances of ~10 s at 16 kHz, ragged), random-init weights.
Prints one JSON line: audio-s/s on the GPU (CUDA events, inputs resident), the fbank kernel against the HBM roofline,
and the oracle (torch fp32, the reference's own arithmetic) on the host cores for one utterance.

    python tools/ppg_bench.py [--batch 32] [--seconds 10] [--iters 10]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seconds", type=float, default=10.0)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--graph", type=int, default=1, help="replay the encoder as one CUDA graph (ASRModel.use_cuda_graph)")
    a = ap.parse_args()
    import ppg_oracle as PO
    from f5e_tts_b200 import _lib
    from f5e_tts_b200.ppg import ASRModel, PPGModelWapper
    dev = torch.device("cuda:0")
    cfg = PO.ConformerConfig(blocks=12)
    sd = PO.make_conformer_state_dict(cfg, seed=7)
    model = ASRModel(cfg.input_dim, num_blocks=cfg.blocks).load_state_dict(sd).to(dev)
    model.use_cuda_graph = bool(a.graph)
    wrap = PPGModelWapper(model=model, device=dev)
    g = torch.Generator().manual_seed(0)
    n = int(a.seconds * 16000)
    lens = torch.randint(int(0.85 * n), n + 1, (a.batch,), generator=g)
    lens[0] = n
    wave = (0.1 * torch.randn(a.batch, n, generator=g)).to(dev)
    lens_d = lens.to(dev)
    audio_s = float(lens.sum()) / 16000.0
    lib = _lib.load()
    for _ in range(3):
        ppg, true_len = wrap.audio_to_ppg(wave, sr=16000, lens=lens_d)
    torch.cuda.synchronize()
    n0 = lib.f5e_launch_count()
    lib.f5e_prof_begin(32)  # F5E_PROF_DSP: the fbank kernel
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        ppg, true_len = wrap.audio_to_ppg(wave, sr=16000, lens=lens_d)
    e1.record()
    torch.cuda.synchronize()
    nrec = lib.f5e_prof_end()
    launches = lib.f5e_launch_count() - n0
    ms = e0.elapsed_time(e1) / a.iters
    cls, ms_, dims = C.c_int32(), C.c_float(), (C.c_int64 * 3)()
    fb_ms, fb_n = 0.0, 0
    for i in range(nrec):
        lib.f5e_prof_read(C.c_int64(i), C.byref(cls), C.byref(ms_), dims)
        if cls.value == 32 and dims[2] == 4:
            fb_ms += ms_.value
            fb_n += 1
    lib.f5e_prof_begin(0)
    frames = int(((lens - 400) // 160 + 1).sum())
    fb_bytes = 4.0 * float(lens.sum()) + 4.0 * 80 * frames  # wave read once + features written once
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", peaks.get("hbm_copy_gbs", 6454.9)) or 6454.9)
    line = dict(metric="PPG extraction (kaldi fbank + 12-block Conformer), source audio seconds per second", unit="audio-s/s",
                value=audio_s / (ms / 1e3), ms_per_batch=ms, batch=a.batch, audio_seconds=audio_s, ppg_frames=int(true_len.sum()),
                cuda_graph=bool(a.graph), host_launches_per_batch=launches // a.iters, dtype="bf16 (fbank fp32)",
                fbank=dict(avg_launch_ms=fb_ms / max(fb_n, 1), bytes=fb_bytes, achieved_gbs=fb_bytes / (fb_ms / max(fb_n, 1) * 1e-3) / 1e9,
                           peak_gbs=hbm, frac=fb_bytes / (fb_ms / max(fb_n, 1) * 1e-3) / 1e9 / hbm))
    if not a.no_cpu:
        torch.set_num_threads(os.cpu_count())
        w0 = wave[0].cpu()
        PO.audio_to_ppg(cfg, sd, w0[:16000])
        t0 = time.perf_counter()
        PO.audio_to_ppg(cfg, sd, w0)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = dict(value=(n / 16000.0) / dt, unit="audio-s/s", cores=os.cpu_count(), kind="port",
                                    sample=f"oracle (torch fp32) on utterance 0 ({n / 16000.0:.1f} s), batch 1")
    print(json.dumps(line))


if __name__ == "__main__":
    main()
