"""Developer tool: timings of the prompt-preparation and long-form rows (one JSON line each).
    python tools/pipeline_bench.py [--prompts 1088] [--chunks 8]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from f5e_tts_b200 import pipeline as P  # noqa: E402
from oracle import f5e_oracle as O  # noqa: E402 (CPU baseline leg only)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--prompts", type=int, default=1088)
    ap.add_argument("--chunks", type=int, default=8)
    ap.add_argument("--cpu-sample", type=int, default=48)
    ap.add_argument("--eval-batches", type=int, default=1000)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    meta = []
    for i in range(a.prompts):  # Seed-TTS-test-en shaped: 16 kHz prompts of 3-6 s, targets giving 3-15 s totals
        secs = 3.0 + 3.0 * torch.rand(1, generator=g).item()
        wave = 0.05 * torch.randn(1, int(secs * 16000), generator=g)
        meta.append((f"u{i}", "reference transcript of the prompt audio.", (wave, 16000),
                     " " + "some generated words " * int(1 + 4 * torch.rand(1, generator=g).item()), None))
    from bench import build_model
    from f5e_tts_b200.vocos import Vocos
    model, _ = build_model(dev)
    model.store_trajectory = False
    voc = Vocos(seed=4321).to(dev)
    kw = dict(infer_batch_size=16 * 1031, num_buckets=200, min_secs=3, max_secs=40)
    P.get_inference_prompt(meta[:64], device=dev, **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = P.get_inference_prompt(meta, device=dev, **kw)
    torch.cuda.synchronize()
    t_gpu = time.perf_counter() - t0
    audio_s = sum(m[2][0].shape[-1] for m in meta) / 16000
    t0 = time.perf_counter()
    O.get_inference_prompt(meta[: a.cpu_sample], **kw)
    t_cpu = (time.perf_counter() - t0) * a.prompts / a.cpu_sample
    print(json.dumps(dict(row="get_inference_prompt", prompts=a.prompts, batches=len(out), prompt_audio_s=round(audio_s, 1),
                          b200_s=round(t_gpu, 3), cpu_port_s_extrapolated=round(t_cpu, 2), cpu_sample=a.cpu_sample,
                          cores=torch.get_num_threads(), note="host wall clock incl. H2D of the prompts and batching logic")))

    # long-form: chunks batched in one solve vs the reference's chunk-by-chunk loop (both through the product)
    ref = (0.05 * torch.randn(1, 5 * 24000, generator=g), 24000)
    ref_text = "a reference transcript of about five seconds of speech, more or less."
    chunks = ["this is a chunk of generated text that is about as long as the chunker allows, one hundred and thirty bytes, "
              "give or take a few."] * a.chunks
    kwi = dict(nfe_step=32, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        return time.perf_counter() - t0, r

    # evaluation set end to end on this GPU (BASELINE configs[3] shape: nfe 16), prompts from above
    from f5e_tts_b200 import evaluate as E
    sub = out[: a.eval_batches]
    E.infer_batches(model, voc, sub[:1], nfe_step=16, seed=0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = E.infer_batches(model, voc, sub, nfe_step=16, seed=0)
    torch.cuda.synchronize()
    t_e = time.perf_counter() - t0
    gen_s = sum(w.shape[-1] for _, w in res) / 24000
    print(json.dumps(dict(row="evaluate.infer_batches", batches=len(sub), utterances=len(res), nfe=16,
                          generated_audio_s=round(gen_s, 1), seconds=round(t_e, 3), audio_s_per_s=round(gen_s / t_e, 1),
                          note="host wall clock incl. per-batch D2H of the waveforms")))

    t_b, rb = timed(lambda: next(P.infer_batch_process(ref, ref_text, chunks, model, voc, **kwi)))
    t_s, _ = timed(lambda: [next(P.infer_batch_process(ref, ref_text, [c], model, voc, **kwi)) for c in chunks])
    print(json.dumps(dict(row="infer_batch_process", chunks=a.chunks, audio_s=round(len(rb[0]) / 24000, 1),
                          batched_s=round(t_b, 3), chunk_by_chunk_s=round(t_s, 3),
                          audio_s_per_s_batched=round(len(rb[0]) / 24000 / t_b, 1))))


if __name__ == "__main__":
    main()
