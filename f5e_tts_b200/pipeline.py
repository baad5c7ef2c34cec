"""Prompt preparation and chunked long-form synthesis around the sampler, on the GPU.

    rms_normalize / resample / prepare_reference   <-> infer/utils_infer.py:436-446, eval/utils_eval.py:119-127
    get_inference_prompt                           <-> eval/utils_eval.py:77-219 (waveforms in, bucketed batches out)
    chunk_text                                     <-> infer/utils_infer.py:70-98
    infer_batch_process                            <-> infer/utils_infer.py:418-564

The reference prepares every prompt on the CPU one file at a time (torchaudio resample + MelSpec) and solves the text
chunks of a long utterance one after the other at batch 1; here the prompts of a whole evaluation set go through three
batched kernels and all chunks are ONE padded solve (`CFM.isolate_samples` makes that equal to the batch-1 solves).
File I/O and the grapheme-to-pinyin conversion (jieba / pypinyin) stay with the caller.
"""
from __future__ import annotations

import ctypes as C
import math
import random
import re

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

TARGET_SAMPLE_RATE, HOP_LENGTH, TARGET_RMS = 24000, 256, 0.1  # infer/utils_infer.py:49-55

_kernel_cache: dict = {}


def sinc_resample_kernel(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """Polyphase filter table of torchaudio.functional.resample (sinc_interp_hann), same float64 arithmetic and the
    same final cast to fp32.  Returns (kern [up, taps] fp32 CPU, width, down, up)."""
    g = math.gcd(int(orig_freq), int(new_freq))
    down, up = int(orig_freq) // g, int(new_freq) // g
    base = min(down, up) * rolloff
    width = math.ceil(lowpass_filter_width * down / base)
    idx = torch.arange(-width, width + down, dtype=torch.float64)[None, :] / down
    t = torch.arange(0, -up, -1, dtype=torch.float64)[:, None] / up + idx
    t = (t * base).clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    kern = torch.where(t == 0, torch.tensor(1.0, dtype=torch.float64), t.sin() / t) * window * (base / down)
    return kern.to(torch.float32).contiguous(), width, down, up


def _lens32(lens, batch, n, dev):
    if lens is None:
        return None
    return torch.as_tensor(lens, device=dev).to(torch.int32).contiguous()


def rms_normalize(wav: torch.Tensor, lens=None, target_rms: float = TARGET_RMS, scale: bool = True):
    """wav fp32 [B, n] (cuda) -> (normalised wav [B, n], rms [B]); rows are zeroed past lens[b]."""
    if wav.device.type != "cuda":
        raise RuntimeError("f5e_tts_b200.pipeline has no CPU path")
    wav = wav.float().contiguous()
    B, n = wav.shape
    lib = _lib.load()
    scratch = torch.empty(max(1, lib.f5e_rms_scratch_floats(B, n)), device=wav.device, dtype=torch.float32)
    rms = torch.empty(B, device=wav.device, dtype=torch.float32)
    out = torch.empty_like(wav) if scale else None
    l32 = _lens32(lens, B, n, wav.device)
    check(lib.f5e_rms_normalize(ptr(wav), C.c_int64(wav.stride(0)), ptr(l32), B, n, C.c_float(target_rms), ptr(scratch),
                                ptr(rms), ptr(out), C.c_int64(n), stream_ptr()), "f5e_rms_normalize")
    return out, rms


def resample(wav: torch.Tensor, orig_freq: int, new_freq: int, lens=None):
    """torchaudio.transforms.Resample(orig_freq, new_freq)(wav) for a padded batch: fp32 [B, n] -> ([B, n'], lens')."""
    if wav.device.type != "cuda":
        raise RuntimeError("f5e_tts_b200.pipeline has no CPU path")
    wav = wav.float().contiguous()
    B, n = wav.shape
    if lens is None:
        lens_t = torch.full((B,), n, dtype=torch.int64)
    else:
        lens_t = torch.as_tensor(lens).to("cpu", torch.int64)
    if int(orig_freq) == int(new_freq):
        return wav, lens_t
    key = (int(orig_freq), int(new_freq), wav.device)
    if key not in _kernel_cache:
        k, width, down, up = sinc_resample_kernel(orig_freq, new_freq)
        _kernel_cache[key] = (k.to(wav.device), width, down, up)
    kern, width, down, up = _kernel_cache[key]
    out_lens = (up * lens_t + down - 1) // down  # ceil(up * length / down)
    out_max = int((up * n + down - 1) // down)
    out = torch.empty(B, out_max, device=wav.device, dtype=torch.float32)
    l32 = lens_t.to(wav.device, torch.int32)
    check(_lib.load().f5e_resample(ptr(wav), C.c_int64(wav.stride(0)), ptr(l32), B, n, ptr(kern), kern.shape[1], down, up,
                                   width, ptr(out), C.c_int64(out_max), out_max, stream_ptr()), "f5e_resample")
    return out, out_lens


def prepare_reference(audio: torch.Tensor, sr: int, target_rms: float = TARGET_RMS,
                      target_sample_rate: int = TARGET_SAMPLE_RATE):
    """infer/utils_infer.py:436-446: channel mean, loudness floor, resample.  audio [C, n] -> (wav [1, n'] cuda, rms)."""
    if audio.shape[0] > 1:
        audio = torch.mean(audio, dim=0, keepdim=True)
    wav, rms = rms_normalize(audio, None, target_rms)
    wav, _ = resample(wav, sr, target_sample_rate)
    return wav, rms[0]


# ---------------------------------------------------------------------------------------------------------------------
# evaluation prompts (eval/utils_eval.py:77-219)
# ---------------------------------------------------------------------------------------------------------------------
def _prepare_references(metainfo, dev, slab, target_sample_rate, n_fft, win_length, n_mel_channels, hop_length,
                        mel_spec_type, target_rms):
    """The audio half of get_inference_prompt / get_vc_inference_prompt (utils_eval.py:118-134, 286-299): loudness floor,
    resampling and MelSpec of the prompts, batched per sample rate, `slab` prompts at a time.
    -> (ref_rms [n], device-resident ref mels [n_mel, frames], ref_mel_len [n])."""
    from .frontend import MelSpec

    if mel_spec_type != "vocos":
        raise NotImplementedError("f5e_tts_b200: only the vocos mel variant is implemented")
    mel_spectrogram = MelSpec(n_fft=n_fft, hop_length=hop_length, win_length=win_length, n_mel_channels=n_mel_channels,
                              target_sample_rate=target_sample_rate, mel_spec_type=mel_spec_type).to(dev)
    ref_rms_all, ref_mel_all, ref_len_all = [None] * len(metainfo), [None] * len(metainfo), [0] * len(metainfo)
    by_sr: dict = {}
    for i, row in enumerate(metainfo):
        wave, sr = row[2]
        wave = wave.reshape(1, -1) if wave.ndim == 1 else wave
        assert wave.shape[0] == 1, "get_inference_prompt expects mono prompts (as torchaudio.load of the eval sets gives)"
        assert wave.shape[-1] > 5000, f"Empty prompt wav: {row[0]}"
        by_sr.setdefault(int(sr), []).append((i, wave.reshape(-1)))
    for sr, items in by_sr.items():
        for s0 in range(0, len(items), slab):
            part = items[s0:s0 + slab]
            lens = torch.tensor([w.numel() for _, w in part])
            batch = torch.zeros(len(part), int(lens.max()), dtype=torch.float32).pin_memory()
            for r, (_, w) in enumerate(part):
                batch[r, : w.numel()] = w
            wav, rms = rms_normalize(batch.to(dev, non_blocking=True), lens, target_rms)
            wav, out_lens = resample(wav, sr, target_sample_rate, lens)
            rms = rms.cpu()
            for r, (i, _) in enumerate(part):
                n = int(out_lens[r])
                ref_rms_all[i] = rms[r]
                ref_len_all[i] = n // hop_length
                ref_mel_all[i] = mel_spectrogram(wav[r:r + 1, :n]).squeeze(0)  # [n_mel, n // hop + 1], device resident
    return ref_rms_all, ref_mel_all, ref_len_all


def _resampled_length(n: int, sr: int, new_sr: int) -> int:
    """Length torchaudio's resampler returns: ceil(new * n / old) on the reduced rates."""
    if int(sr) == int(new_sr):
        return n
    g = math.gcd(int(sr), int(new_sr))
    return -(-(int(new_sr) // g) * n // (int(sr) // g))


def get_inference_prompt(metainfo, speed=1.0, text_converter=None, target_sample_rate=TARGET_SAMPLE_RATE, n_fft=1024,
                         win_length=1024, n_mel_channels=100, hop_length=HOP_LENGTH, mel_spec_type="vocos",
                         target_rms=TARGET_RMS, use_truth_duration=False, infer_batch_size=1, num_buckets=200, min_secs=3,
                         max_secs=40, device="cuda", slab=128):
    """Same bucketing / batching / shuffling as the reference, same return tuples
    (utts, ref_rms_list, padded ref mels [B, T, n_mel], ref_mel_lens, total_mel_lens, final_text_list) per batch.
    metainfo rows: (utt, prompt_text, (prompt_wave [1, n] or [n], sr), gt_text, gt) where `gt` is (wave, sr) when
    use_truth_duration else ignored - waveforms instead of file names: decoding audio files stays with the caller.
    text_converter: the tokenizer front end (`convert_char_to_pinyin` in the reference); default keeps the raw strings.
    The loudness floor, resampling and MelSpec of `slab` prompts at a time run as batched kernels on `device`."""
    dev = torch.device(device)
    metainfo = list(metainfo)
    ref_rms_all, ref_mel_all, ref_len_all = _prepare_references(metainfo, dev, slab, target_sample_rate, n_fft, win_length,
                                                               n_mel_channels, hop_length, mel_spec_type, target_rms)
    # ---- text, durations, buckets: host integer logic, identical to the reference ----
    def fields(i, row):
        prompt_text = row[1]
        if len(prompt_text[-1].encode("utf-8")) == 1:
            prompt_text = prompt_text + " "
        text = [prompt_text + row[3]]
        return prompt_text, (ref_rms_all[i], ref_mel_all[i], ref_len_all[i], text_converter(text) if text_converter is not None else text)

    def emit(utts, total_mel_lens, cols):
        ref_rms_list, ref_mels, ref_mel_lens, texts = cols
        return (utts, ref_rms_list, _pad_ref_mels(ref_mels), ref_mel_lens, total_mel_lens, [t for tl in texts for t in tl])

    return _bucket_batches(metainfo, ref_len_all, fields, emit, speed, target_sample_rate, hop_length, use_truth_duration,
                           infer_batch_size, num_buckets, min_secs, max_secs)


def _pad_ref_mels(ref_mels):
    T = max(m.shape[-1] for m in ref_mels)
    return torch.stack([torch.nn.functional.pad(m, (0, T - m.shape[-1]), value=0) for m in ref_mels]).permute(0, 2, 1)


def _bucket_batches(metainfo, ref_len_all, fields, emit, speed, target_sample_rate, hop_length, use_truth_duration,
                    infer_batch_size, num_buckets, min_secs, max_secs):
    """The duration estimate / bucket / flush / shuffle loop that get_inference_prompt (eval/utils_eval.py:131-219) and
    get_vc_inference_prompt (:322-397) share: host integer logic whose bucket contents and order must equal the reference's.
    fields(i, row) -> (prompt_text used for the duration estimate, per-utterance columns); emit(utts, total_mel_lens, columns)
    -> the tuple of one flushed batch."""
    prompts_all = []
    min_tokens = min_secs * target_sample_rate // hop_length
    max_tokens = max_secs * target_sample_rate // hop_length
    batch_accum = [0] * num_buckets
    utts, total_mel_lens = [[] for _ in range(num_buckets)], [[] for _ in range(num_buckets)]
    cols = [None] * num_buckets

    def flush(b):
        prompts_all.append(emit(utts[b], total_mel_lens[b], cols[b]))
        batch_accum[b], utts[b], total_mel_lens[b], cols[b] = 0, [], [], None

    for i, row in enumerate(metainfo):
        utt, gt_text = row[0], row[3]
        prompt_text, item = fields(i, row)
        ref_mel_len = ref_len_all[i]
        if use_truth_duration:
            gt_wave, gt_sr = row[4]
            n_gt = _resampled_length(gt_wave.shape[-1], gt_sr, target_sample_rate)
            total_mel_len = ref_mel_len + int(n_gt / hop_length / speed)
        else:
            ref_text_len = len(prompt_text.encode("utf-8"))
            gen_text_len = len(gt_text.encode("utf-8"))
            total_mel_len = ref_mel_len + int(ref_mel_len / ref_text_len * gen_text_len / speed)
        assert infer_batch_size > 0, "infer_batch_size should be greater than 0."
        assert min_tokens <= total_mel_len <= max_tokens, (
            f"Audio {utt} has duration {total_mel_len * hop_length // target_sample_rate}s out of range [{min_secs}, {max_secs}].")
        b = math.floor((total_mel_len - min_tokens) / (max_tokens - min_tokens + 1) * num_buckets)
        if cols[b] is None:
            cols[b] = tuple([] for _ in item)
        utts[b].append(utt)
        total_mel_lens[b].append(total_mel_len)
        for c, v in zip(cols[b], item):
            c.append(v)
        batch_accum[b] += total_mel_len
        if batch_accum[b] >= infer_batch_size:
            flush(b)
    for b, frames in enumerate(batch_accum):
        if frames > 0:
            flush(b)
    random.seed(666)  # "not only leave easy work for last workers" (utils_eval.py:214-216)
    random.shuffle(prompts_all)
    return prompts_all


def get_vc_inference_prompt(metainfo, ppg_model, speed=1.0, target_sample_rate=TARGET_SAMPLE_RATE, n_fft=1024, win_length=1024,
                            n_mel_channels=100, hop_length=HOP_LENGTH, mel_spec_type="vocos", target_rms=TARGET_RMS,
                            use_truth_duration=False, infer_batch_size=1, num_buckets=200, min_secs=3, max_secs=40,
                            device="cuda", slab=64):
    """eval/utils_eval.py:245-397 with waveforms instead of file names: rows (utt, prompt_text, (prompt_wave, sr), gt_text,
    (gt_wave, gt_sr)).  Same return tuples (utts, ref_rms_list, padded ref mels [B, T, n_mel], ref_mel_lens, total_mel_lens,
    padded PPG [B, T_ppg, dim]), same bucketing and shuffle.  As in the reference (:301-320) the PPG is extracted from the
    SOURCE audio concatenated with itself (the "prompt" half is loaded from gt_wav there too), resampled to 16 kHz, multi-channel
    sources averaged.  `slab` sources at a time go through the resampler, kaldi fbank and the Conformer as one ragged batch."""
    dev = torch.device(device)
    metainfo = list(metainfo)
    ref_rms_all, ref_mel_all, ref_len_all = _prepare_references(metainfo, dev, slab, target_sample_rate, n_fft, win_length,
                                                               n_mel_channels, hop_length, mel_spec_type, target_rms)
    # ---- PPG of the doubled source, batched per sample rate ----
    ppg_all, ppg_len_all = [None] * len(metainfo), [0] * len(metainfo)
    by_sr: dict = {}
    for i, row in enumerate(metainfo):
        wave, sr = row[4]
        wave = wave.reshape(1, -1) if wave.ndim == 1 else wave
        by_sr.setdefault((int(sr), wave.shape[0]), []).append((i, wave))
    for (sr, ch), items in by_sr.items():
        for s0 in range(0, len(items), slab):
            part = items[s0:s0 + slab]
            lens = torch.tensor([w.shape[-1] for _, w in part])
            n_max = int(lens.max())
            batch = torch.zeros(len(part) * ch, n_max, dtype=torch.float32).pin_memory()
            for r, (_, w) in enumerate(part):
                batch[r * ch:(r + 1) * ch, : w.shape[-1]] = w
            wav, out_lens = resample(batch.to(dev, non_blocking=True), sr, 16000, lens.repeat_interleave(ch))
            if ch > 1:  # resample every channel, then average (:301-308)
                wav, out_lens = wav.view(len(part), ch, -1).mean(dim=1), out_lens[::ch]
            n16 = out_lens.to(dev)
            m = wav.shape[1]
            idx = torch.arange(2 * m, device=dev).unsqueeze(0)
            src = torch.where(idx < n16[:, None], idx, idx - n16[:, None]).clamp(0, m - 1)
            full = torch.gather(wav, 1, src) * (idx < 2 * n16[:, None])  # [source | source], ragged
            ppg, true_len = ppg_model.audio_to_ppg(full, 16000, lens=2 * n16)
            true_len = true_len.cpu()
            for r, (i, _) in enumerate(part):
                t = (int(1 + (2 * int(out_lens[r]) - 400) // 160) - 1) // 2  # frames the batch-1 call of the reference returns
                ppg_all[i], ppg_len_all[i] = ppg[r, :t], int(true_len[r])
    # ---- durations, buckets: host integer logic, identical to the reference ----
    def fields(i, row):  # the VC loop estimates the duration on the RAW prompt text (no space appended, :350-357)
        return row[1], (ref_rms_all[i], ref_mel_all[i], ref_len_all[i], ppg_all[i], ppg_len_all[i])

    def emit(utts, total_mel_lens, cols):
        ref_rms_list, ref_mels, ref_mel_lens, ppg_list, ppg_lens = cols
        P = max(ppg_lens)  # padded_ppg_batch (:226-242): truncate / pad every PPG to the longest ppg_length of the batch
        pp = torch.stack([torch.nn.functional.pad(p[:P], (0, 0, 0, P - min(p.shape[0], P)), value=0) for p in ppg_list])
        return (utts, ref_rms_list, _pad_ref_mels(ref_mels), ref_mel_lens, total_mel_lens, pp)

    return _bucket_batches(metainfo, ref_len_all, fields, emit, speed, target_sample_rate, hop_length, use_truth_duration,
                           infer_batch_size, num_buckets, min_secs, max_secs)


# ---------------------------------------------------------------------------------------------------------------------
# long-form synthesis (infer/utils_infer.py:70-98, 418-564)
# ---------------------------------------------------------------------------------------------------------------------
_SENTENCE_END = re.compile(r"(?<=[;:,.!?])\s+|(?<=[；：，。！？])")


def chunk_text(text: str, max_chars: int = 135) -> list:
    """Greedy packing of sentences (split after ASCII punctuation + whitespace, or after full-width punctuation) into
    chunks of at most `max_chars` UTF-8 bytes; a sentence ending in a single-byte character is followed by a space."""
    nbytes = lambda s: len(s.encode("utf-8"))  # noqa: E731
    out, cur = [], ""
    for sent in _SENTENCE_END.split(text):
        piece = sent + " " if sent and nbytes(sent[-1]) == 1 else sent
        if nbytes(cur) + nbytes(sent) > max_chars:
            if cur:
                out.append(cur.strip())
            cur = piece
        else:
            cur += piece
    if cur:
        out.append(cur.strip())
    return out


def crossfade_chunks(waves: torch.Tensor, lens, cross_fade_duration: float = 0.15, sample_rate: int = TARGET_SAMPLE_RATE,
                     gain=None) -> torch.Tensor:
    """The numpy chain of utils_infer.py:525-556 on the device: waves fp32 [B, n_max] (chunk b valid up to lens[b]) ->
    float64 [n_total].  gain = (rms, target_rms) applies `wave * rms / target_rms` to every chunk first (fp32)."""
    lens = [int(v) for v in lens]
    waves = waves.float().contiguous()
    acc = torch.empty(max(1, sum(lens)), device=waves.device, dtype=torch.float64)
    cf = int(cross_fade_duration * sample_rate) if cross_fade_duration > 0 else 0
    gn, gd = (float(gain[0]), float(gain[1])) if gain is not None else (1.0, 1.0)
    n_acc = C.c_int64(0)
    lib = _lib.load()
    for b, n in enumerate(lens):
        check(lib.f5e_crossfade_append(ptr(acc), n_acc, C.c_void_p(waves[b].data_ptr()), n, cf, C.c_float(gn), C.c_float(gd),
                                       int(gain is not None), C.byref(n_acc), stream_ptr()), "f5e_crossfade_append")
    return acc[: n_acc.value]


def infer_batch_process(ref_audio, ref_text, gen_text_batches, model_obj, vocoder, mel_spec_type="vocos", progress=None,
                        target_rms=TARGET_RMS, cross_fade_duration=0.15, nfe_step=32, cfg_strength=2.0,
                        sway_sampling_coef=-1, speed=1, fix_duration=None, device=None, streaming=False, chunk_size=2048,
                        text_converter=None, seed=None):
    """Generator with the reference's contract (utils_infer.py:418-564): yields (final_wave float64 numpy, sample rate,
    combined spectrogram [n_mel, T]) once, or (chunk, sample rate) pieces when streaming.  All text chunks are solved as
    ONE padded batch (the reference loops over them at batch 1) with `isolate_samples`, decoded by one batched Vocos call
    and cross-faded on the device; only the final wave crosses PCIe."""
    if mel_spec_type != "vocos":
        raise NotImplementedError("f5e_tts_b200: only the vocos mel variant is implemented")
    audio, sr = ref_audio
    dev = torch.device(device) if device is not None else model_obj.device
    audio, rms = prepare_reference(audio.to(dev), sr, target_rms, TARGET_SAMPLE_RATE)
    if len(ref_text[-1].encode("utf-8")) == 1:
        ref_text = ref_text + " "
    gen_text_batches = list(gen_text_batches)
    if not gen_text_batches:
        yield None, TARGET_SAMPLE_RATE, None
        return
    ref_audio_len = audio.shape[-1] // HOP_LENGTH
    texts, durations = [], []
    for gen_text in gen_text_batches:
        local_speed = 0.3 if len(gen_text.encode("utf-8")) < 10 else speed
        text_list = [ref_text + gen_text]
        texts.extend(text_converter(text_list) if text_converter is not None else text_list)
        if fix_duration is not None:
            durations.append(int(fix_duration * TARGET_SAMPLE_RATE / HOP_LENGTH))
        else:
            ref_text_len, gen_text_len = len(ref_text.encode("utf-8")), len(gen_text.encode("utf-8"))
            durations.append(ref_audio_len + int(ref_audio_len / ref_text_len * gen_text_len / local_speed))
    B = len(texts)
    cond = model_obj.mel_spec(audio).permute(0, 2, 1)  # once; CFM.sample would compute the same mel for every chunk
    keep = model_obj.isolate_samples
    model_obj.isolate_samples = True
    try:
        generated, _ = model_obj.sample(cond=cond.expand(B, -1, -1), text=texts,
                                        duration=torch.tensor(durations, device=dev, dtype=torch.long), steps=nfe_step,
                                        cfg_strength=cfg_strength, sway_sampling_coef=sway_sampling_coef, seed=seed)
    finally:
        model_obj.isolate_samples = keep
    del _
    # CFM.sample raises short durations to max(text, prompt) + 1 (cfm.py:405-407); a batch-1 call returns that many frames
    lens = cond.shape[1]
    if isinstance(texts[0], str) and getattr(model_obj, "vocab_char_map", None) is None:
        n_text = [len(t.encode("utf-8")) for t in texts]  # list_str_to_tensor: UTF-8 bytes are the tokens
    else:
        n_text = [len(t) for t in texts]
    if getattr(model_obj, "vocab_char_map", None) is not None and (model_obj.use_align_loss or model_obj.use_cross_mask):
        n_text = [2 * n + 1 for n in n_text]  # interspersed blanks (cfm.py:130-131)
    eff = [max(max(nt, lens) + 1, d) for nt, d in zip(n_text, durations)]
    start = torch.full((B,), ref_audio_len, device=dev, dtype=torch.long)
    end = torch.tensor(eff, device=dev, dtype=torch.long)
    waves = vocoder.decode_varlen(generated.float(), start, end, max_frames=max(eff) - ref_audio_len)
    wav_lens = [(e - ref_audio_len - 1) * HOP_LENGTH for e in eff]
    low = bool(rms < target_rms)
    gain = (float(rms), float(target_rms)) if low else None
    if streaming:
        for b in range(B):
            w = waves[b, : wav_lens[b]]
            if low:
                w = w * rms / target_rms
            w = w.cpu().numpy()
            for j in range(0, len(w), chunk_size):
                yield w[j: j + chunk_size], TARGET_SAMPLE_RATE
        return
    if cross_fade_duration <= 0:
        parts = [waves[b, : wav_lens[b]] for b in range(B)]
        final = torch.cat([p * rms / target_rms if low else p for p in parts])
    else:
        final = crossfade_chunks(waves, wav_lens, cross_fade_duration, TARGET_SAMPLE_RATE, gain)
    spec = torch.cat([generated[b, ref_audio_len: eff[b]].float().permute(1, 0) for b in range(B)], dim=1)
    yield final.cpu().numpy(), TARGET_SAMPLE_RATE, spec.cpu().numpy()
