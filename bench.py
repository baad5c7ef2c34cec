#!/usr/bin/env python
"""Headline benchmark: audio-seconds generated per second, F5TTS_v1_Base, NFE 32, CFG 2, sway -1, bf16, + Vocos decode.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # the B200 path (libf5e_b200.so)
    python bench.py --impl reference [...]                         # the reference algorithm on the host cores (oracle port)

A "step" is one pass of the hot path over one batch: CFM.sample (32 Euler steps x (cond + uncond) DiT evaluations, fused
CFG + Euler) on 16 utterances of ~10 s, then the Vocos decode of every generated span (BASELINE.json configs[1]).
Under torchrun every rank runs its own batch (per-utterance sharding, no data-path collective inside the solve); the e2e leg
adds the NCCL gather of the waveforms to rank 0.  One JSON line is printed by rank 0.

Timed regions of the B200 arm, in order (each bracketed by barrier + synchronize, CUDA events, max over ranks):
  1. `value`    inputs resident in HBM, in-stream profiler OFF;
  2. `e2e`      raw prompt waves in pinned host memory -> H2D -> MelSpec -> CFM.sample -> Vocos -> (NCCL gather) -> D2H into
                pinned memory; the gather + D2H of step i run on a side stream under the solve of step i + 1;
  3. roofline   ONE more step with every launch of the GEMM / attention / norm classes bracketed by events
                (f5e_prof_begin) -> `roofline` (DiT-block GEMMs), `roofline_attention`, `roofline_hbm`;
  4. extras     `gpu_baseline` (the path restated in eager PyTorch bf16 on the same GPU, 2 NFE steps, with the reference's
                expanded bool mask and with flash SDPA), `extra.latency` (batch 1), `extra.eval_set` (BASELINE configs[3]
                shape sharded over the ranks: strong scaling), `cpu_baseline` (rank 0, N = 1 only).
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
    ap.add_argument("--graph", type=int, default=1, help="CFM.use_cuda_graph: the Euler loop of a solve as ONE cached CUDA graph "
                    "launch (same kernels; 1.5 %% faster at config 2). The per-kernel roofline step always runs on the stream")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=15.0)
    ap.add_argument("--no-extras", action="store_true", help="skip gpu_baseline / latency / eval_set")
    ap.add_argument("--eval-utterances", type=int, default=1088)
    ap.add_argument("--pack", type=int, default=-1, help="CFM.pack_valid_rows: -1 default (exact: padded rows kept, attention sub-block skipped on them), 0 padded layout, 1 drop padded rows")
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
def _reference_model(sd_fp32):
    """The UNMODIFIED reference (oracle/_ref, copied from /root/reference by oracle/make_ref.py at build time) with the bench
    weights, on the CPU in fp32; None when oracle/_ref has not been made (then the oracle port is timed instead)."""
    import torch

    try:
        from oracle.make_ref import import_reference

        CFM, DiT = import_reference()
    except Exception:
        return None
    if CFM is None:
        return None
    dit = DiT(**ARCH)
    missing, unexpected = dit.load_state_dict(sd_fp32, strict=False)
    assert not unexpected, unexpected
    model = CFM(transformer=dit, mel_spec_kwargs=dict()).eval()
    return model.to(torch.float32)


def cpu_arm(sd_fp32, w, nfe, budget_s):
    """Times `CFM.sample`'s per-step work for ONE utterance of the workload on the host cores and extrapolates linearly
    to the full solve (+ the Vocos decode, timed in full on the oracle's restatement: the `vocos` package is not installed).
    The sampler is the reference's own code when oracle/_ref exists (kind "reference"), else the oracle port (kind "port").
    Returns audio-s/s and a description of the sample."""
    import torch

    from oracle import f5e_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    i = 0
    ref_len, dur = int(w["ref_len"][i]), int(w["dur"][i])
    mel = O.mel_spectrogram(w["ref_wav"][i: i + 1, : ref_len * HOP]).permute(0, 2, 1)[:, :ref_len]
    text = w["text"][i: i + 1]
    text = text[:, : int((text != -1).sum())]
    ref_model = _reference_model(sd_fp32)
    if ref_model is not None:
        kind = "reference"

        def solve(k):
            with torch.no_grad():
                ref_model.sample(cond=mel, text=text, ppg=None, duration=dur, lens=torch.tensor([ref_len]), steps=k, cfg_strength=2.0,
                                 sway_sampling_coef=-1.0, seed=0)
    else:
        kind = "port"
        dit = O.OracleDiT(O.DiTConfig(), sd_fp32, torch.float32, "cpu")

        def solve(k):
            O.cfm_sample(dit, mel, text, dur, lens=torch.tensor([ref_len]), steps=k, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)
    steps_done, t_used = 0, 0.0
    k = 1
    while True:  # k Euler steps per probe until the budget is used; the per-step cost is what is extrapolated
        t0 = time.perf_counter()
        solve(k)
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
    what = ("the unmodified reference (oracle/_ref: f5_tts.model.CFM.sample, torch fp32)" if kind == "reference"
            else "oracle port (torch fp32)")
    return dict(value=audio / (total + voc_s), unit=UNIT, cores=cores, kind=kind, steps_timed=steps_done,
                sample=f"{what} on {cores} host threads: utterance 0 of the workload ({dur} frames, {ref_len} ref), "
                       f"{steps_done} of {nfe} Euler steps timed ({t_used:.1f} s) and extrapolated linearly, Vocos decode "
                       f"(oracle restatement; the vocos package is absent) timed in full"), per_step


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
    vals, per, timed_steps = [], [], 0
    budget = max(2.0, min(a.cpu_budget_s, 120.0 / max(1, a.steps + a.warmup)))
    r = None
    for s in range(a.warmup + a.steps):
        r, per_step = cpu_arm(sd, w, a.nfe, budget)
        if s >= a.warmup:
            vals.append(r["value"])
            per.append(per_step)
            timed_steps = r["steps_timed"]
    v = statistics.mean(vals)
    r["value"] = v
    line = dict(metric=METRIC, value=v, unit=UNIT, impl="reference", n_gpus=a.gpus, steps=a.steps, warmup=a.warmup,
                ms_per_step=1e3 * statistics.mean(per) * a.nfe, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic",
                config=dict(workload_config(a, w), cuda_graph=bool(a.graph)),  # same keys / values as the B200 arm's config
                extrapolated=True, steps_timed=timed_steps,
                extrapolation="ONE utterance of the batch at batch 1, steps_timed of nfe Euler steps per bench step, scaled "
                              "linearly to nfe steps; audio-s/s of that utterance (per-utterance cost is what shards)",
                cpu_baseline=r,
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
MUFU_LANES_PER_CLK_SM = 16.0   # MUFU.EX2 results per clock per SM (measured: tools/ubench/tmem_bw.cu, 15.8)
N_SMS = 148


def profile_step(lib, step_fn, dur, B, peak_tf, peak_gbs, sm_mhz):
    """One more step with every GEMM / attention / norm launch bracketed by an event pair on the launching stream
    -> the three roofline blocks of the JSON line.  Algorithmic work only: valid frames, no padding credited."""
    import ctypes as C

    import torch

    lib.f5e_prof_begin(1 | 2 | 8)  # F5E_PROF_GEMM | ATTN | NORM
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    step_fn()
    e1.record()
    torch.cuda.synchronize()
    ms_total = e0.elapsed_time(e1)
    nrec = lib.f5e_prof_end()
    rows_valid = 2.0 * float(dur.sum())
    rows_padded = 2 * B * int(dur.max())
    sum_n2 = 2.0 * float((dur.double() ** 2).sum())
    g_ms = g_fl = a_ms = r_ms = r_by = 0.0
    g_n = a_n = r_n = 0
    cls, ms_, dims = C.c_int32(), C.c_float(), (C.c_int64 * 3)()
    for i in range(nrec):
        lib.f5e_prof_read(C.c_int64(i), C.byref(cls), C.byref(ms_), dims)
        d0, d1, d2 = dims[0], dims[1], dims[2]
        if cls.value == 1:
            M, N, K = d0, d1, d2
            if (N, K - LN_FOLD_K) in ((3072, 1024), (2048, 1024)):
                K -= LN_FOLD_K  # the mean-correction block is overhead, not algorithmic work
            if M in (rows_padded, int(rows_valid)) and (N, K) in DIT_GEMM_NK:  # padded or packed rows
                g_ms += ms_.value
                g_fl += 2.0 * rows_valid * N * K
                g_n += 1
        elif cls.value == 2 and d0 == 2 * B:  # the DiT blocks' attention (not the text / PPG embedders')
            a_ms += ms_.value
            a_n += 1
        elif cls.value == 8 and d2 == 1 and d1 == 1024 and d0 in (rows_padded, int(rows_valid)):  # row_stats of the DiT blocks
            r_ms += ms_.value
            r_by += rows_valid * (d1 * 2 + 4 + 64 * 2)  # bf16 row in, fp32 rstd + bf16 [64] mean-correction row out
            r_n += 1
    lib.f5e_prof_begin(0)
    out = {}
    ach = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else None
    out["gemm"] = dict(achieved=ach, launches=g_n, avg_launch_ms=(g_ms / g_n if g_n else None),
                       share_of_step=(g_ms / ms_total if ms_total else None), frac=(ach / peak_tf if ach else None))
    if a_n:
        fl = 4.0 * 1024 * sum_n2  # QK^T + PV over the valid keys / queries of the doubled batch, 16 heads x 64
        avg = a_ms / a_n
        tf = fl / (avg * 1e-3) / 1e12
        floor_ms = 16.0 * sum_n2 / (MUFU_LANES_PER_CLK_SM * N_SMS * sm_mhz * 1e6) * 1e3  # one ex2 per (query, key, head)
        out["attention"] = dict(bound="mufu+tensor", kernel="attention_kernel (DiT blocks)", flops_per_launch=fl, achieved=tf,
                                peak=peak_tf, unit="TFLOP/s", frac=tf / peak_tf, launches=a_n, avg_launch_ms=avg,
                                share_of_step=a_ms / ms_total, mufu_floor_ms=floor_ms, frac_of_mufu_floor=floor_ms / avg,
                                mufu_floor="16 heads x sum N^2 exponentials / (16 lanes/clk/SM x 148 SMs x the SM clock sampled "
                                           "during the timed region)")
    if r_n:
        avg = r_ms / r_n
        gbs = r_by / r_n / (avg * 1e-3) / 1e9
        out["hbm"] = dict(bound="hbm", kernel="row_stats_kernel (LayerNorm statistics of the folded adaLN path)",
                          bytes_per_launch=r_by / r_n, achieved=gbs, peak=peak_gbs, unit="GB/s", frac=gbs / peak_gbs,
                          launches=r_n, avg_launch_ms=avg, share_of_step=r_ms / ms_total)
    return out, ms_total


def gpu_baseline(sd_fp32, w, dev, nfe_steps=2):
    """The kernel to beat (BASELINE.md B-GPU-ref): the path restated in eager PyTorch bf16 (cuBLAS, SDPA, cuDNN convs) on the
    SAME GPU and workload, `nfe_steps` Euler steps (cond + uncond evaluation each).  Once with the reference's expanded bool
    mask (modules.py:483-492: the math / mem-efficient SDPA backends) and once with flash SDPA (per sample, valid keys only)."""
    import torch

    from oracle import f5e_oracle as O

    cfg = O.DiTConfig(**ARCH)
    o16 = O.OracleDiT(cfg, sd_fp32, torch.bfloat16, dev)
    ref_len, dur, text = w["ref_len"].to(dev), w["dur"].to(dev), w["text"].to(dev)
    mel = O.mel_spectrogram(w["ref_wav"].to(dev)).permute(0, 2, 1)[:, : int(ref_len.max())].contiguous()
    kw = dict(steps=nfe_steps, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)
    res = {}
    for name, backend in (("bool_mask", "mask"), ("flash", "flash")):
        o16.attn_backend = backend
        fn = lambda: O.cfm_sample(o16, mel, text, dur, lens=ref_len, **kw)  # noqa: E731
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(2):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 2 / nfe_steps
    return dict(kind="oracle port, eager PyTorch bf16 (cuBLAS / SDPA / cuDNN), same GPU, same workload",
                nfe_steps_timed=nfe_steps, ms_per_nfe_step_bool_mask=res["bool_mask"], ms_per_nfe_step_flash=res["flash"])


def latency_leg(model, vocoder, dev, nfe):
    """BASELINE.md "C1 shape on GPU": ONE ~10 s utterance (940 frames, 315 of them prompt), nfe Euler steps, CFG 2 - what
    infer_cli.py runs.  Steady state: the CUDA graph of the Euler loop is cached by shape after the first solve."""
    import torch

    g = torch.Generator().manual_seed(7)
    cond = torch.randn(1, 315, 100, generator=g).to(dev)
    text = torch.randint(1, 2545, (1, 150), generator=g).to(dev)
    keep = model.use_cuda_graph
    out = {}
    for graph in (0, 1):
        model.use_cuda_graph = bool(graph)
        fn = lambda: model.sample(cond, text, None, 940, steps=nfe, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)  # noqa: E731
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out["graph" if graph else "stream"] = e0.elapsed_time(e1) / 3
    model.use_cuda_graph = keep
    best = min(out.values())
    return dict(workload="batch 1 x 940 frames (315 prompt), nfe %d, cfg 2, mel only" % nfe, ms_per_solve_stream=out["stream"],
                ms_per_solve_graph=out["graph"], ms_per_nfe_step=best / nfe, floor_ms_per_nfe_step=0.62,
                audio_s_per_s=(940 - 315) * HOP / SR / (best / 1e3))


def eval_set_leg(model, vocoder, dev, rank, world, n_utt, nfe=16, batch_frames=32 * 1031, num_buckets=60):
    """BASELINE configs[3]: a Seed-TTS-test-en shaped synthetic set (n_utt utterances of 3-15 s, NFE 16) - a FIXED total
    sharded over the ranks (strong scaling) through evaluate.split_batches ("dynamic": a shared counter), then the ONE
    exchange of the path: lengths + one padded NCCL gather of the waveforms to rank 0."""
    import torch
    import torch.distributed as dist

    from f5e_tts_b200 import dist as fdist
    from f5e_tts_b200 import evaluate as E
    from f5e_tts_b200 import pipeline as P

    g = torch.Generator().manual_seed(0)  # the same list on every rank, like the reference
    meta = []
    for i in range(n_utt):
        secs = 3.0 + 3.0 * torch.rand(1, generator=g).item()
        wave = 0.05 * torch.randn(1, int(secs * 16000), generator=g)
        meta.append((f"u{i}", "reference transcript of the prompt audio.", (wave, 16000),
                     " " + "some generated words " * int(1 + 4 * torch.rand(1, generator=g).item()), None))
    prompts = P.get_inference_prompt(meta, device=dev, infer_batch_size=batch_frames, num_buckets=num_buckets, min_secs=3, max_secs=40)
    E.infer_batches(model, vocoder, prompts[:1], nfe_step=nfe, seed=0)  # warm-up
    if world > 1:  # NCCL builds its peer channels lazily and the allocator has not seen buffers of this size yet: warm both
        for n in (1, 32 << 20):
            fdist.gather_varlen(torch.ones(1, device=dev, dtype=torch.int64) * n, torch.zeros(n, device=dev), dst=0)
        dist.barrier()
    # rank 0's pinned staging buffer for the whole set (a long-lived buffer in a real job: page-locking 600 MB is not free)
    n_all = sum((t - r - 1) * HOP for b in prompts for t, r in zip(b[4], b[3]))
    host = torch.empty(n_all, dtype=torch.float32, pin_memory=True) if rank == 0 else None
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()  # a common start: the job's wall time is the slowest rank's time from here to the end of the gather
    t0 = t_start = time.perf_counter()
    # the shard's waves stay on the device; ONE exchange (lengths + a padded NCCL gather to rank 0) and ONE copy to the host
    res = E.infer_batches(model, vocoder, prompts, nfe_step=nfe, seed=0, rank=rank, world=world, split="dynamic",
                          keep_on_device=True)  # batches drawn from one shared counter, dearest first (evaluate._DynamicBatches)
    lens = torch.tensor([x.shape[-1] for _, x in res], device=dev)
    payload = torch.cat([x.reshape(-1) for _, x in res]) if res else torch.zeros(0, device=dev)
    torch.cuda.synchronize()
    t_mine = time.perf_counter() - t0
    t0 = time.perf_counter()
    gathered = fdist.gather_varlen(lens, payload, dst=0) if world > 1 else (lens, payload)
    if rank == 0:
        assert gathered[1].numel() == n_all
        host.copy_(gathered[1], non_blocking=True)
    torch.cuda.synchronize()
    t_gather = time.perf_counter() - t_start  # this rank's WHOLE time (solve, wait for the slowest rank, gather, D2H)
    t = torch.tensor([t_mine, t_gather, float(len(res)), float(lens.sum())], device=dev, dtype=torch.float64)
    tmax, tsum = t.clone(), t.clone()
    busy = [t_mine]
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        allb = torch.zeros(world, device=dev, dtype=torch.float64)
        allb[rank] = t_mine
        dist.all_reduce(allb)
        busy = allb.tolist()
    if rank != 0:
        return None
    audio = float(tsum[3]) / SR
    assert int(tsum[2]) == n_utt and (world == 1 or int(gathered[0].numel()) == n_utt)
    mean_busy = sum(busy) / len(busy)
    return dict(workload=f"BASELINE configs[3] shape: {n_utt} synthetic utterances (3-6 s prompts, 3-15 s total), {len(prompts)} "
                         f"bucketed batches, nfe {nfe}, cfg 2, Vocos decode; fixed total sharded over the ranks, the waves gathered over "
                         f"NCCL to rank 0 and copied to pinned host memory once (gather_ms)",
                scaling="strong", n_gpus=world, generated_audio_s=audio, audio_s_per_s=audio / float(tmax[1]), wall_s=float(tmax[1]),
                solve_s_max_over_ranks=float(tmax[0]), per_rank_busy_s=[round(b, 3) for b in busy],
                lpt_imbalance=max(busy) / mean_busy if mean_busy > 0 else None, split="dynamic (shared counter)", gather_ms=1e3 * (float(tmax[1]) - float(tmax[0])),  # what the exchange + D2H add behind the slowest rank's solve
                gather_bytes=int(tsum[3]) * 4, timing="host wall clock around synchronised regions (the rank's batches; then gather + "
                                                      "D2H), max over ranks")


def run_b200(a, rank, local_rank, world):
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
    if a.pack >= 0:
        model.pack_valid_rows = bool(a.pack)
    from f5e_tts_b200.vocos import Vocos

    vocoder = Vocos(seed=4321).to(dev)
    w = make_workload(a.batch, a.seconds, rank)  # each rank owns its own utterances
    B = a.batch
    ref_len, dur, text = w["ref_len"], w["dur"], w["text"]
    kw = dict(steps=a.nfe, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)

    # device-resident inputs of region 1: the prompt mels (MelSpec kernel), token ids, lengths
    mel = model.mel_spec(w["ref_wav"].to(dev)).permute(0, 2, 1)[:, : int(ref_len.max())].contiguous()  # [B, Tref, 100] fp32
    d_cond, d_text, d_ref, d_dur = mel, text.to(dev), ref_len.to(dev), dur.to(dev)
    # host buffers of region 2: RAW prompt waves (CFM.sample takes them, cfm.py:371-374), token ids, lengths; result buffer
    h_wave, h_text = w["ref_wav"].pin_memory(), text.pin_memory()
    h_ref, h_dur = ref_len.pin_memory(), dur.pin_memory()
    gen_len = (dur - ref_len)
    wav_len = (gen_len - 1) * HOP
    l_loc = torch.tensor([int(wav_len.max())], device=dev, dtype=torch.int64)
    if world > 1:
        dist.all_reduce(l_loc, op=dist.ReduceOp.MAX)
    l_max = int(l_loc)
    h_out = torch.empty(world * B if rank == 0 else 1, l_max, dtype=torch.float32).pin_memory()
    side = torch.cuda.Stream(device=dev)
    gather_ms = []

    def step_device():
        out, _ = model.sample(d_cond, d_text, None, d_dur, lens=d_ref, **kw)
        return vocoder.decode_varlen(out, d_ref, d_dur)  # fp32 [B, max (gen-1)*256], zero padded

    def step_e2e():
        wave = h_wave.to(dev, non_blocking=True)
        txt = h_text.to(dev, non_blocking=True)
        rl = h_ref.to(dev, non_blocking=True)
        du = h_dur.to(dev, non_blocking=True)
        out, _ = model.sample(wave, txt, None, du, lens=rl, **kw)   # MelSpec of the prompts runs inside
        wav = vocoder.decode_varlen(out, rl, du)
        # delivery (gather to rank 0 over NVLink, then D2H into pinned memory) on a side stream: it overlaps the next solve
        main = torch.cuda.current_stream()
        side.wait_stream(main)
        with torch.cuda.stream(side):
            wav.record_stream(side)
            if wav.shape[1] != l_max:
                wav = torch.nn.functional.pad(wav, (0, l_max - wav.shape[1]))
            if world > 1:
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                g0.record(side)
                if rank == 0:
                    allw = torch.empty(world, B, l_max, device=dev, dtype=torch.float32)
                    dist.gather(wav.contiguous(), list(allw.unbind(0)), dst=0)
                    wav = allw.view(world * B, l_max)
                else:
                    dist.gather(wav.contiguous(), None, dst=0)
                g1.record(side)
                gather_ms.append((g0, g1))
            if rank == 0:
                h_out.copy_(wav, non_blocking=True)

    def timed(fn, n, drain=None):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        if drain is not None:
            torch.cuda.current_stream().wait_stream(drain)  # the last step's delivery is inside the timed region
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
    # --- timed region 1: inputs resident in HBM, profiler off
    clocks = ClockSampler(local_rank)
    n0 = lib.f5e_launch_count()
    torch.cuda.profiler.start()  # cudaProfilerStart: `ncu --profile-from-start off` captures the timed region only
    ms_total = timed(step_device, a.steps)
    torch.cuda.profiler.stop()
    launches = lib.f5e_launch_count() - n0
    clk = clocks.stop()
    ms_step = ms_total / a.steps
    audio = audio_seconds(w)
    tot_audio = torch.tensor([audio], device=dev, dtype=torch.float64)  # ranks own different lengths: sum the audio
    if world > 1:
        dist.all_reduce(tot_audio)
    value = float(tot_audio) / (ms_step / 1e3)

    # --- timed region 2: end to end through the public API with HOST buffers
    for _ in range(2):
        step_e2e()
    torch.cuda.synchronize()
    gather_ms.clear()
    ms_e2e = timed(step_e2e, a.steps, drain=side) / a.steps
    e2e_value = float(tot_audio) / (ms_e2e / 1e3)
    h2d = h_wave.numel() * 4 + h_text.numel() * 8 + h_ref.numel() * 8 + h_dur.numel() * 8
    d2h = h_out.numel() * 4 if rank == 0 else 0
    g_ms = [x.elapsed_time(y) for x, y in gather_ms] if gather_ms else []

    # --- region 3: one profiled step -> rooflines (rank 0's numbers are reported)
    peak_tf, peak_gbs, peak_src = measured_peaks()
    model.use_cuda_graph = False  # event pairs around single launches need the stream path
    step_device()
    roof, ms_prof = profile_step(lib, step_device, dur, B, peak_tf, peak_gbs, (clk or {}).get("sm_mhz", 1965.0))
    model.use_cuda_graph = bool(a.graph)

    extra = {}
    if not a.no_extras:
        ev = eval_set_leg(model, vocoder, dev, rank, world, a.eval_utterances)
        if rank == 0:
            extra["eval_set"] = ev
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp))
    step_flops = 2.0 * sum(dit_flops(float(n)) for n in dur)
    gm = roof["gemm"]
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=a.steps, warmup=a.warmup, ms_per_step=ms_step,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic",
                config=dict(workload_config(a, w), cuda_graph=bool(a.graph)),
                ms_per_nfe_step=ms_step / a.nfe,
                model_tflops=step_flops * a.nfe / (ms_step * 1e-3) / 1e12,
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h),
                         ms_per_step=ms_e2e, inputs="raw prompt waves (MelSpec inside the timed region), token ids, lengths",
                         delivery="gather + D2H of step i on a side stream under the solve of step i+1; the last one drained "
                                  "inside the timed region",
                         gather_ms_per_step=(statistics.mean(g_ms) if g_ms else 0.0)),
                gpu_launches=int(launches),
                pack_valid_rows=("drop" if model.last_aux["dropped_padded_rows"] else
                                 "split (exact)" if model.last_aux["packed_rows"] else "off"),
                roofline=dict(bound="tensor", kernel="gemm_bf16_kernel (DiT block qkv/out/ff1/ff2)", achieved=gm["achieved"],
                              peak=peak_tf, unit="TFLOP/s", frac=gm["frac"], traffic=traffic.get("gemm_bf16_kernel"),
                              peak_source=peak_src, launches=gm["launches"], avg_launch_ms=gm["avg_launch_ms"],
                              share_of_step=gm["share_of_step"],
                              measured="in-stream event pairs around every launch of ONE extra step (profiler off in the "
                                       "headline region); %.1f ms with the events vs %.1f ms without" % (ms_prof, ms_step)),
                clocks=clk)
    if "attention" in roof:
        line["roofline_attention"] = dict(roof["attention"], traffic=traffic.get("attention_kernel"))
    if "hbm" in roof:
        line["roofline_hbm"] = dict(roof["hbm"], traffic=traffic.get("row_stats_kernel"), peak_source=peak_src)
    if not a.no_extras:
        try:
            line["gpu_baseline"] = gpu_baseline(sd_fp32, w, dev)
        except Exception as e:  # the baseline arm must never take the headline down with it
            line["gpu_baseline"] = dict(error=repr(e)[:200])
        try:
            extra["latency"] = latency_leg(model, vocoder, dev, a.nfe)
        except Exception as e:
            extra["latency"] = dict(error=repr(e)[:200])
    if extra:
        line["extra"] = extra
    if not a.no_cpu_baseline and world == 1:
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
