#!/usr/bin/env python
"""Headline benchmark: audio-seconds generated per second, F5TTS_v1_Base, NFE 32, CFG 2, sway -1, bf16, + Vocos decode.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # the B200 path (libf5e_b200.so)
    python bench.py --impl reference [...]                         # the reference algorithm on the host cores (oracle port)

A "step" is one pass of the hot path over one batch: CFM.sample (32 Euler steps x (cond + uncond) DiT evaluations, fused
CFG + Euler) on 16 utterances of ~10 s, then the Vocos decode of every generated span (BASELINE.json configs[1]).
Under torchrun every rank runs its own batch (per-utterance sharding, no data-path collective inside the solve); the e2e leg
adds the NCCL gather of the waveforms to rank 0.  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "audio-sec/sec F5TTS_v1_Base NFE32 CFG2 (generated audio only)"
UNIT = "audio-s/s"
HOP, SR = 256, 24000
ARCH = dict(dim=1024, depth=22, heads=16, ff_mult=2, text_dim=512, conv_layers=4, text_num_embeds=2545, mel_dim=100)
DIT_GEMM_NK = {(3072, 1024), (1024, 1024), (2048, 1024), (1024, 2048)}  # qkv / out / ff1 / ff2 of a DiTBlock
LN_FOLD_K = 64  # qkv / ff1 with LayerNorm folded in run with K = dim + 64 (one extra block, one UMMA_K step of it issued)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=16, help="utterances per GPU per step (config 2: 16)")
    ap.add_argument("--nfe", type=int, default=32)
    ap.add_argument("--seconds", type=float, default=10.0, help="mean utterance length (config 5: 30)")
    ap.add_argument("--graph", type=int, default=0, help="capture the Euler loop in one CUDA graph per solve")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=15.0)
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------------------------
# workload (SURVEY 8(d) "C2 inputs"): synthetic, seeded, identical on every run
# ----------------------------------------------------------------------------------------------------------------------
def make_workload(batch: int, seconds: float, seed: int):
    import torch

    g = torch.Generator().manual_seed(seed)
    scale = seconds / 10.0
    ref_len = torch.randint(int(235 * scale), int(421 * scale), (batch,), generator=g)
    dur = torch.randint(int(850 * scale), int(1031 * scale), (batch,), generator=g)
    tlen = torch.randint(int(100 * scale), int(201 * scale), (batch,), generator=g)
    text = torch.full((batch, int(tlen.max())), -1, dtype=torch.long)
    for i in range(batch):
        text[i, : tlen[i]] = torch.randint(1, 2545, (int(tlen[i]),), generator=g)
    wav = 0.1 * torch.randn(batch, int(ref_len.max()) * HOP, generator=g)
    return dict(ref_len=ref_len, dur=dur, text=text, ref_wav=wav)


def audio_seconds(w):
    return float((w["dur"] - w["ref_len"]).sum()) * HOP / SR


def dit_flops(n: float) -> float:
    """Algorithmic FLOPs of one DiT evaluation of one sample with n valid frames."""
    return n * (378.89e6 + 90112.0 * n)


def build_model(device):
    import torch

    from f5e_tts_b200.model import CFM, DiT

    torch.manual_seed(1234)
    dit = DiT(**ARCH)
    with torch.no_grad():  # the reference zero-initialises these; zeros make the model output identically 0
        for blk in dit.transformer_blocks:
            blk.attn_norm.linear.weight.normal_(0, 0.02)
            blk.attn_norm.linear.bias.normal_(0, 0.02)
        for m in (dit.norm_out.linear, dit.proj_out):
            m.weight.normal_(0, 0.02)
            m.bias.normal_(0, 0.02)
    model = CFM(dit)
    sd_fp32 = {k: v.detach().clone() for k, v in dit.state_dict().items()}
    return model.to(device).to(torch.bfloat16), sd_fp32


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return None
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, r[3:7]):
                if v.strip().lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return None
        return dict(sm_mhz=statistics.median(sm), sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", d.get("bf16_tflops")), d.get("hbm_gbs"), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm (oracle port, torch fp32) on the host cores
# ----------------------------------------------------------------------------------------------------------------------
def cpu_arm(sd_fp32, w, nfe, budget_s):
    """Times `CFM.sample`'s per-step work for ONE utterance of the workload on the host cores and extrapolates linearly
    to the full solve (+ the oracle Vocos decode, timed in full).  Returns audio-s/s and a description of the sample."""
    import torch

    from oracle import f5e_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = O.DiTConfig()
    dit = O.OracleDiT(cfg, sd_fp32, torch.float32, "cpu")
    i = 0
    ref_len, dur = int(w["ref_len"][i]), int(w["dur"][i])
    mel = O.mel_spectrogram(w["ref_wav"][i: i + 1, : ref_len * HOP]).permute(0, 2, 1)[:, :ref_len]
    text = w["text"][i: i + 1]
    text = text[:, : int((text != -1).sum())]
    steps_done, t_used = 0, 0.0
    k = 1
    while True:  # k Euler steps per probe until the budget is used; the per-step cost is what is extrapolated
        t0 = time.perf_counter()
        O.cfm_sample(dit, mel, text, dur, lens=torch.tensor([ref_len]), steps=k, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)
        t_used += time.perf_counter() - t0
        steps_done += k
        if t_used >= budget_s or steps_done >= nfe:
            break
    per_step = t_used / steps_done
    total = per_step * nfe
    voc_s = 0.0
    if hasattr(O, "vocos_decode"):
        vsd = O.make_vocos_state_dict(4321)
        m = torch.randn(1, 100, dur - ref_len)
        t0 = time.perf_counter()
        O.vocos_decode(vsd, m)
        voc_s = time.perf_counter() - t0
    audio = (dur - ref_len) * HOP / SR
    return dict(value=audio / (total + voc_s), unit=UNIT, cores=cores, kind="port",
                sample=f"oracle port (torch fp32) on {cores} host threads: utterance 0 of the workload ({dur} frames, {ref_len} ref), "
                       f"{steps_done} of {nfe} Euler steps timed ({t_used:.1f} s) and extrapolated linearly, Vocos decode timed in full"), per_step


def reference_weights():
    """The same seeded weights as the B200 arm, on the CPU (the drop-in DiT is used as a parameter container only)."""
    import torch

    from f5e_tts_b200.model import DiT

    torch.manual_seed(1234)
    dit = DiT(**ARCH)
    with torch.no_grad():
        for blk in dit.transformer_blocks:
            blk.attn_norm.linear.weight.normal_(0, 0.02)
            blk.attn_norm.linear.bias.normal_(0, 0.02)
        for m in (dit.norm_out.linear, dit.proj_out):
            m.weight.normal_(0, 0.02)
            m.bias.normal_(0, 0.02)
    return {k: v.detach() for k, v in dit.state_dict().items()}


def run_reference(a, rank):
    """--impl reference: the reference algorithm (oracle port) on the host cores; rank 0 only."""
    if rank != 0:
        return
    w = make_workload(a.batch, a.seconds, 0)
    sd = reference_weights()
    vals, per = [], []
    budget = max(2.0, min(a.cpu_budget_s, 120.0 / max(1, a.steps + a.warmup)))
    r = None
    for s in range(a.warmup + a.steps):
        r, per_step = cpu_arm(sd, w, a.nfe, budget)
        if s >= a.warmup:
            vals.append(r["value"])
            per.append(per_step)
    v = statistics.mean(vals)
    r["value"] = v
    line = dict(metric=METRIC, value=v, unit=UNIT, impl="reference", n_gpus=a.gpus, steps=a.steps, warmup=a.warmup,
                ms_per_step=1e3 * statistics.mean(per) * a.nfe, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic",
                config=workload_config(a, w), cpu_baseline=r,
                e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def workload_config(a, w):
    return dict(workload=f"BASELINE configs[1]: F5TTS_v1_Base (dim 1024, depth 22, 16 heads) random-init + random-init Vocos mel-24khz, "
                         f"batch {a.batch} x ~{a.seconds:g} s utterances per GPU, Euler nfe={a.nfe}, cfg=2, sway=-1",
                batch_per_gpu=a.batch, frames_total=int(w["dur"].sum()), frames_generated=int((w["dur"] - w["ref_len"]).sum()),
                n_pad=int(w["dur"].max()), nfe=a.nfe, cfg_strength=2.0, sway=-1.0, sharding="per-utterance",
                l2="activations+weights per step exceed the 126 MB L2 (inputs larger than L2; no explicit flush)")


# ----------------------------------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------------------------------
def run_b200(a, rank, local_rank, world):
    import ctypes as C

    import torch
    import torch.distributed as dist

    from f5e_tts_b200 import _lib

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    _lib.require_device()
    lib = _lib.load()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model, sd_fp32 = build_model(dev)
    model.use_cuda_graph = bool(a.graph)
    model.store_trajectory = False  # every caller of the reference discards the trajectory
    from f5e_tts_b200.vocos import Vocos

    vocoder = Vocos(seed=4321).to(dev)
    w = make_workload(a.batch, a.seconds, rank)  # each rank owns its own utterances
    B = a.batch
    ref_len, dur, text = w["ref_len"], w["dur"], w["text"]
    kw = dict(steps=a.nfe, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)

    # reference prompt mels through the MelSpec kernel (once, outside the timed region: it is input preparation)
    mel = model.mel_spec(w["ref_wav"].to(dev)).permute(0, 2, 1)[:, : int(ref_len.max())].contiguous()  # [B, Tref, 100] fp32
    d_cond, d_text, d_ref, d_dur = mel, text.to(dev), ref_len.to(dev), dur.to(dev)
    h_cond, h_text = mel.cpu().pin_memory(), text.pin_memory()
    h_ref, h_dur = ref_len.pin_memory(), dur.pin_memory()
    gen_len = (dur - ref_len)
    wav_len = (gen_len - 1) * HOP
    h_wav = torch.empty(B, int(wav_len.max()), dtype=torch.float32).pin_memory()

    def step_device():
        out, _ = model.sample(d_cond, d_text, None, d_dur, lens=d_ref, **kw)
        return vocoder.decode_varlen(out, d_ref, d_dur)  # fp32 [B, max (gen-1)*256], zero padded

    def step_e2e():
        cond = h_cond.to(dev, non_blocking=True)
        txt = h_text.to(dev, non_blocking=True)
        rl = h_ref.to(dev, non_blocking=True)
        du = h_dur.to(dev, non_blocking=True)
        out, _ = model.sample(cond, txt, None, du, lens=rl, **kw)
        wav = vocoder.decode_varlen(out, rl, du)
        if world > 1:  # the one exchange of the sharded eval: gather every rank's waveforms on rank 0 over NVLink
            from f5e_tts_b200 import dist as fdist

            wav = fdist.gather_padded(wav, dst=0)
        if rank == 0:
            tgt = h_wav if world == 1 else None
            if tgt is not None:
                tgt.copy_(wav, non_blocking=True)
            else:
                wav.cpu()
        torch.cuda.current_stream().synchronize()

    def timed(fn, n):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for _ in range(a.warmup):
        step_device()
    # --- timed region 1: inputs resident in HBM; GEMM launches bracketed by events for the roofline leg
    clocks = ClockSampler(local_rank)
    n0 = lib.f5e_launch_count()
    lib.f5e_prof_begin(1)  # F5E_PROF_GEMM
    torch.cuda.profiler.start()  # cudaProfilerStart: `ncu --profile-from-start off` captures the timed region only
    ms_total = timed(step_device, a.steps)
    torch.cuda.profiler.stop()
    nrec = lib.f5e_prof_end()
    launches = lib.f5e_launch_count() - n0
    clk = clocks.stop()
    # per-launch records -> achieved TFLOP/s of the DiT-block GEMM kernel
    rows_valid = 2.0 * float(dur.sum())
    n_pad = int(dur.max())
    rows_padded = 2 * B * n_pad
    g_ms, g_fl, g_n = 0.0, 0.0, 0
    cls, ms_, dims = C.c_int32(), C.c_float(), (C.c_int64 * 3)()
    for i in range(nrec):
        lib.f5e_prof_read(C.c_int64(i), C.byref(cls), C.byref(ms_), dims)
        M, N, K = dims[0], dims[1], dims[2]
        if (N, K - LN_FOLD_K) in ((3072, 1024), (2048, 1024)):
            K -= LN_FOLD_K  # the mean-correction block is overhead, not algorithmic work
        if cls.value == 1 and M in (rows_padded, int(rows_valid)) and (N, K) in DIT_GEMM_NK:  # padded or packed rows
            g_ms += ms_.value
            g_fl += 2.0 * rows_valid * N * K  # algorithmic: valid rows only, padding not credited
            g_n += 1
    lib.f5e_prof_begin(0)
    ms_step = ms_total / a.steps
    audio = audio_seconds(w)
    # the other ranks' workloads have different lengths: sum the audio over ranks
    tot_audio = torch.tensor([audio], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot_audio)
    value = float(tot_audio) / (ms_step / 1e3)

    # --- timed region 2: end to end through the public API with HOST buffers
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, a.steps) / a.steps
    e2e_value = float(tot_audio) / (ms_e2e / 1e3)
    h2d = h_cond.numel() * 4 + h_text.numel() * 8 + h_ref.numel() * 8 + h_dur.numel() * 8
    d2h = h_wav.numel() * 4 * (world if world > 1 else 1)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak_tf, peak_gbs, peak_src = measured_peaks()
    ach = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else None
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("gemm_bf16_kernel")
    step_flops = 2.0 * sum(dit_flops(float(n)) for n in dur)
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=a.steps, warmup=a.warmup, ms_per_step=ms_step,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic",
                config=dict(workload_config(a, w), cuda_graph=bool(a.graph)),
                ms_per_nfe_step=ms_step / a.nfe,
                model_tflops=step_flops * a.nfe / (ms_step * 1e-3) / 1e12,
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h),
                         ms_per_step=ms_e2e),
                gpu_launches=int(launches),
                roofline=dict(bound="tensor", kernel="gemm_bf16_kernel (DiT block qkv/out/ff1/ff2)", achieved=ach, peak=peak_tf,
                              unit="TFLOP/s", frac=(ach / peak_tf if ach else None), traffic=traffic, peak_source=peak_src,
                              launches=g_n, avg_launch_ms=(g_ms / g_n if g_n else None),
                              share_of_step=(g_ms / ms_total if ms_total else None)),
                clocks=clk)
    if not a.no_cpu_baseline:
        cb, _ = cpu_arm(sd_fp32, make_workload(a.batch, a.seconds, 0), a.nfe, a.cpu_budget_s)
        line["cpu_baseline"] = cb
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if a.impl == "reference":
        run_reference(a, rank)
    else:
        run_b200(a, rank, local_rank, world)


if __name__ == "__main__":
    main()
