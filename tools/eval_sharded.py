"""BASELINE configs[3]: a Seed-TTS-test-en shaped synthetic evaluation set (1088 utterances of 3-15 s), NFE 16, sharded by
batch across the ranks of one box; one JSON line from rank 0.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/eval_sharded.py"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from bench import build_model  # noqa: E402
from f5e_tts_b200 import dist as fdist  # noqa: E402
from f5e_tts_b200 import evaluate as E  # noqa: E402
from f5e_tts_b200 import pipeline as P  # noqa: E402
from f5e_tts_b200.vocos import Vocos  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utterances", type=int, default=1088)
    ap.add_argument("--nfe", type=int, default=16)
    ap.add_argument("--split", default="lpt")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model, _ = build_model(dev)
    model.store_trajectory = False
    voc = Vocos(seed=4321).to(dev)
    g = torch.Generator().manual_seed(0)  # the same set on every rank, like the reference (every rank reads the same list)
    meta = []
    for i in range(a.utterances):
        secs = 3.0 + 3.0 * torch.rand(1, generator=g).item()
        wave = 0.05 * torch.randn(1, int(secs * 16000), generator=g)
        meta.append((f"u{i}", "reference transcript of the prompt audio.", (wave, 16000),
                     " " + "some generated words " * int(1 + 4 * torch.rand(1, generator=g).item()), None))
    t0 = time.perf_counter()
    prompts = P.get_inference_prompt(meta, device=dev, infer_batch_size=16 * 1031, num_buckets=200, min_secs=3, max_secs=40)
    torch.cuda.synchronize()
    t_prep = time.perf_counter() - t0
    E.infer_batches(model, voc, prompts[:1], nfe_step=a.nfe, seed=0)  # warm-up
    if world > 1:  # NCCL sets its peer connections up lazily: warm the gather path before timing it
        fdist.gather_varlen(torch.ones(1, device=dev, dtype=torch.int64), torch.zeros(1, device=dev), dst=0)
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = E.infer_batches(model, voc, prompts, nfe_step=a.nfe, seed=0, rank=rank, world=world, split=a.split)
    torch.cuda.synchronize()
    t_mine = time.perf_counter() - t0
    # the one exchange: lengths, then ONE padded gather of the waveforms to rank 0 over NVLink
    lens = torch.tensor([w.shape[-1] for _, w in res], device=dev)
    payload = torch.cat([w.reshape(-1) for _, w in res]).to(dev) if res else torch.zeros(0, device=dev)
    t0 = time.perf_counter()
    gathered = fdist.gather_varlen(lens, payload, dst=0) if world > 1 else (lens, payload)
    torch.cuda.synchronize()
    t_gather = time.perf_counter() - t0
    t = torch.tensor([t_mine, t_gather, float(len(res)), float(lens.sum())], device=dev, dtype=torch.float64)
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    if rank == 0:
        audio = float(tsum[3]) / 24000
        assert int(tsum[2]) == a.utterances
        if world > 1:
            assert int(gathered[0].numel()) == a.utterances and int(gathered[1].numel()) == int(tsum[3])
        print(json.dumps(dict(row="eval set sharded", n_gpus=world, utterances=a.utterances, batches=len(prompts), nfe=a.nfe,
                              split=a.split, generated_audio_s=round(audio, 1), solve_s_max_over_ranks=round(float(tmax[0]), 3),
                              solve_s_mean=round(float(tsum[0]) / world, 3), gather_s=round(float(tmax[1]), 4),
                              prompt_prep_s=round(t_prep, 3), audio_s_per_s=round(audio / (float(tmax[0]) + float(tmax[1])), 1))))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
