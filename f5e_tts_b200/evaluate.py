"""Batched evaluation inference, sharded by batch across ranks: the loop of eval/eval_infer_batch.py:184-220
(and its _tts / _vc siblings) over the prompts `pipeline.get_inference_prompt` produced.

The reference splits the shuffled batch list contiguously between processes (`accelerator.split_between_processes`),
decodes every utterance with its own Vocos call, copies each wave to the host and writes a file.  Here a rank solves its
batches (LPT-balanced on the modelled cost by default, `split="reference"` keeps the contiguous split), decodes a whole
batch with one `Vocos.decode_varlen`, applies the loudness gain on the device and hands (utt, wave) to a sink - writing
files (torchaudio.save in the reference) stays with the caller.  The only exchange is the optional final gather."""
from __future__ import annotations

import torch

from . import dist as fdist

TARGET_SAMPLE_RATE, HOP_LENGTH = 24000, 256


def split_batches(prompts_all, rank: int, world: int, split: str = "lpt") -> list:
    """Indices of the batches rank `rank` solves.  "reference": contiguous chunks like accelerate's
    split_between_processes (the first len % world ranks get one more); "lpt": longest-processing-time on the modelled
    cost sum(a N + b N^2) of each batch (dist.utterance_cost), deterministic, identical on every rank; "dynamic": an
    iterator that draws from a counter shared by the ranks (_DynamicBatches; needs an initialised process group)."""
    n = len(prompts_all)
    if world <= 1:
        return list(range(n))
    if split == "reference":
        base, extra = divmod(n, world)
        start = rank * base + min(rank, extra)
        return list(range(start, start + base + (1 if rank < extra else 0)))
    if split == "dynamic":
        return _DynamicBatches(prompts_all)
    if split != "lpt":
        raise ValueError(f"split must be 'lpt', 'dynamic' or 'reference', not {split!r}")
    cost = [sum(fdist.utterance_cost(d) for d in p[4]) for p in prompts_all]
    order = sorted(range(n), key=lambda i: (-cost[i], i))
    load, mine = [0.0] * world, []
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += cost[i]
        if r == rank:
            mine.append(i)
    return sorted(mine)


_dynamic_calls = 0


class _DynamicBatches:
    """split="dynamic": the ranks draw batch indices from ONE shared counter (an atomic add on the process group's
    rendezvous store, ~0.1 ms per batch), most expensive batch first.  A static plan cannot know that the GPUs of a box differ by
    a few per cent or that a rank was delayed; with the counter every rank stays busy until the list is empty.  Which rank solves
    a batch does not change its result.  Every rank must iterate the same sequence of infer_batches calls (the counter's key
    is the call number)."""

    def __init__(self, prompts_all):
        import torch.distributed as dist
        global _dynamic_calls
        cost = [sum(fdist.utterance_cost(d) for d in p[4]) for p in prompts_all]
        self.order = sorted(range(len(prompts_all)), key=lambda i: (-cost[i], i))
        self.store = dist.distributed_c10d._get_default_store()
        self.key = f"f5e_tts_b200/dynamic_batches/{_dynamic_calls}"
        _dynamic_calls += 1

    def __iter__(self):
        while True:
            k = self.store.add(self.key, 1) - 1
            if k >= len(self.order):
                return
            yield self.order[k]


@torch.no_grad()
def infer_batches(model, vocoder, prompts_all, *, nfe_step=32, cfg_strength=2.0, sway_sampling_coef=-1.0,
                  no_ref_audio=False, seed=None, target_rms=0.1, rank=0, world=1, split="lpt", sink=None, device=None,
                  sampler="sample", sampler_kwargs=None, keep_on_device=False):
    """Solve this rank's batches.  Returns [(utt, wave fp32 [1, n] on the CPU)] (or whatever `sink(utt, wave, sr)`
    returned, when a sink is given).  sampler: "sample" (eval_infer_batch.py), "sample_tts" (eval_infer_batch_tts.py:
    alpha_spk / alpha_txt in sampler_kwargs instead of cfg_strength), "sample_vc" (eval_infer_batch_vc.py: batches from
    pipeline.get_vc_inference_prompt carry the padded PPG where the text list is; alpha_spk / alpha_ppg in sampler_kwargs).
    keep_on_device: the waves stay on the GPU (no D2H copy and no host synchronisation per batch): for callers that gather the
    shards over NCCL first and copy to the host once (dist.gather_varlen), instead of the reference's per-utterance file
    writes from CPU tensors (eval_infer_batch.py:208-218)."""
    dev = torch.device(device) if device is not None else model.device
    out = []
    for bi in split_batches(prompts_all, rank, world, split):
        utts, ref_rms_list, ref_mels, ref_mel_lens, total_mel_lens, final_text_list = prompts_all[bi]  # [5]: PPG for sample_vc
        ref_mels = ref_mels.to(dev)
        ref_len = torch.tensor(ref_mel_lens, dtype=torch.long, device=dev)
        total = torch.tensor(total_mel_lens, dtype=torch.long, device=dev)
        if sampler == "sample":
            generated, _ = model.sample(cond=ref_mels, text=final_text_list, duration=total, lens=ref_len, steps=nfe_step,
                                        cfg_strength=cfg_strength, sway_sampling_coef=sway_sampling_coef,
                                        no_ref_audio=no_ref_audio, seed=seed)
        elif sampler == "sample_tts":
            generated, _ = model.sample_tts(cond=ref_mels, text=final_text_list, duration=total, lens=ref_len, steps=nfe_step,
                                            sway_sampling_coef=sway_sampling_coef, no_ref_audio=no_ref_audio, seed=seed,
                                            **(sampler_kwargs or {}))
        elif sampler == "sample_vc":  # eval_infer_batch_vc.py:215-225 over pipeline.get_vc_inference_prompt batches
            generated, _ = model.sample_vc(cond=ref_mels, ppg=final_text_list.to(dev), duration=total, lens=ref_len, steps=nfe_step,
                                           sway_sampling_coef=sway_sampling_coef, no_ref_audio=no_ref_audio, seed=seed,
                                           **(sampler_kwargs or {}))
        else:
            raise ValueError(f"unknown sampler {sampler!r}")
        n_frames = [t - r for t, r in zip(total_mel_lens, ref_mel_lens)]
        waves = vocoder.decode_varlen(generated, ref_len, total, max_frames=max(n_frames))  # [B, (max - 1) * hop]
        rms = torch.stack([torch.as_tensor(r, dtype=torch.float32) for r in ref_rms_list]).to(dev)
        # `if ref_rms < target_rms: wave = wave * ref_rms / target_rms` (eval_infer_batch.py:214-215), per row
        scaled = waves * rms[:, None] / target_rms
        waves = torch.where((rms < target_rms)[:, None], scaled, waves)
        if not keep_on_device:
            waves = waves.cpu()
        for i, utt in enumerate(utts):
            w = waves[i: i + 1, : (n_frames[i] - 1) * HOP_LENGTH]
            out.append(sink(utt, w, TARGET_SAMPLE_RATE) if sink is not None else (utt, w))
    return out
