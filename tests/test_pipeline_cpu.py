"""CPU: host logic of the prompt-preparation / long-form pipeline and the oracle restatements, against fixtures made
from the reference's own code (tests/golden/resample.pt: torchaudio; chunk_text.pt: infer/utils_infer.py:chunk_text)."""
import os

import numpy as np
import torch

from oracle import f5e_oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


def test_oracle_resample_matches_torchaudio_fixture():
    for c in torch.load(os.path.join(G, "resample.pt"), weights_only=False):
        y = O.resample(c["x"], c["orig"], c["new"])
        assert y.shape == c["y"].shape
        assert (y - c["y"]).abs().max().item() < 5e-6, (c["orig"], c["new"])


def test_filter_table_is_the_oracle_table():
    """The product builds its polyphase table on the host: same float64 arithmetic as the restated torchaudio kernel."""
    from f5e_tts_b200.pipeline import sinc_resample_kernel
    for orig, new in [(16000, 24000), (44100, 24000), (48000, 24000), (22050, 24000)]:
        kern, width, down, up = sinc_resample_kernel(orig, new)
        assert kern.shape == (up, 2 * width + down)
        # a unit impulse through the oracle's conv1d reads the table back (reversed tap order)
        x = torch.zeros(1, 4 * down + 2 * width + 1)
        pos = 2 * down + width
        x[0, pos] = 1.0
        y = O.resample(x, orig, new)
        for p in range(up):
            for q in range(1, 4):
                k = pos + width - q * down
                if 0 <= k < kern.shape[1]:
                    assert y[0, q * up + p].item() == kern[p, k].item()


def test_chunk_text_matches_reference_fixture():
    from f5e_tts_b200.pipeline import chunk_text
    for row in torch.load(os.path.join(G, "chunk_text.pt"), weights_only=False):
        assert chunk_text(row["text"], max_chars=row["max_chars"]) == row["chunks"], row["text"][:30]


def test_crossfade_chain_properties():
    rng = np.random.default_rng(0)
    waves = [rng.standard_normal(n).astype(np.float32) for n in (9000, 5000, 2000, 12000)]
    out = O.crossfade_chain(waves, 0.15, 24000)
    cf = [min(3600, a, b) for a, b in [(9000, 5000), (9000 + 5000 - 3600, 2000), (9000 + 5000 - 3600, 12000)]]
    assert out.dtype == np.float64 and len(out) == sum(map(len, waves)) - 3600 - 2000 - 3600
    assert np.array_equal(out[:9000 - 3600], waves[0][:9000 - 3600].astype(np.float64))
    assert np.array_equal(out[-(12000 - 3600):], waves[3][3600:].astype(np.float64))
    assert np.array_equal(O.crossfade_chain(waves, 0.0), np.concatenate(waves))
    del cf


def test_split_batches_covers_every_batch_once():
    from f5e_tts_b200.evaluate import split_batches
    rng = np.random.default_rng(1)
    prompts = [(None, None, None, None, list(rng.integers(300, 1400, size=int(rng.integers(1, 17)))), None) for _ in range(37)]
    for split in ("lpt", "reference"):
        for world in (1, 2, 4, 8):
            parts = [split_batches(prompts, r, world, split) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(37))
            assert parts == [split_batches(prompts, r, world, split) for r in range(world)]  # deterministic
    # accelerate's split_between_processes: contiguous, first ranks one longer
    assert split_batches(prompts, 0, 4, "reference") == list(range(0, 10))
    assert split_batches(prompts, 3, 4, "reference") == list(range(28, 37))
    # LPT balances the modelled cost
    from f5e_tts_b200.dist import utterance_cost
    cost = [sum(utterance_cost(d) for d in p[4]) for p in prompts]
    loads = [sum(cost[i] for i in split_batches(prompts, r, 4, "lpt")) for r in range(4)]
    assert max(loads) / (sum(loads) / 4) < 1.1
