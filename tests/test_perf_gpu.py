"""The "kernel to beat" of BASELINE.md (row B-GPU-ref): eager PyTorch (cuBLAS / flash SDPA / cuDNN) running the path's
restatement in bf16 on the SAME B200, against the product, on bench.py's workload (configs[1]: 16 utterances of ~10 s).
The unmodified reference cannot travel to the GPU box (pure Python + absent third-party packages), so the oracle - pinned
against it in the build container - stands in for it.  A measurement printed by a test, with one weak assertion (the
product must not be slower than eager PyTorch); the bench numbers proper come from bench.py."""
import os
import sys

import pytest
import torch

ROOT = os.path.join(os.path.dirname(__file__), "..")
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]

pytestmark = pytest.mark.gpu


def _time(fn, n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def test_product_against_eager_pytorch_on_the_same_gpu(cuda_dev):
    import bench
    import f5e_oracle as O
    nfe = 8  # per-step cost does not depend on the step count; 8 keeps the eager arm short
    model, sd = bench.build_model(cuda_dev)
    model.store_trajectory = False
    w = bench.make_workload(16, 10.0, 0)
    ref_len, dur, text = w["ref_len"].to(cuda_dev), w["dur"].to(cuda_dev), w["text"].to(cuda_dev)
    mel = model.mel_spec(w["ref_wav"].to(cuda_dev)).permute(0, 2, 1)[:, : int(ref_len.max())].contiguous()
    kw = dict(steps=nfe, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=0)
    cfg = O.DiTConfig(**bench.ARCH)
    o16 = O.OracleDiT(cfg, sd, torch.bfloat16, cuda_dev)
    product = lambda: model.sample(mel, text, None, dur, lens=ref_len, **kw)  # noqa: E731
    eager = lambda: O.cfm_sample(o16, mel, text, dur, lens=ref_len, **kw)  # noqa: E731
    for _ in range(2):
        product(), eager()
    ms_p, ms_e = _time(product, 3) / nfe, _time(eager, 3) / nfe
    print(f"\nconfigs[1] shape, per NFE step (cond + uncond evaluation, CFG, Euler): product {ms_p:.2f} ms, "
          f"eager PyTorch bf16 on the same B200 {ms_e:.2f} ms  ->  {ms_e / ms_p:.2f}x")
    assert ms_p < ms_e
