"""CPU tests: the C-ABI library loads and exports every declared symbol; host-side sharding; the world_size-2 gather on gloo."""
import ctypes
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from f5e_tts_b200 import _lib, build
    lib = build.build()
    h = ctypes.CDLL(str(lib))
    syms = _lib.exported_symbols()
    assert len(syms) >= 24 and "f5e_ode_solve" in syms and "f5e_vocos_decode" in syms and "f5e_vq_quantize" in syms
    for s in syms:
        assert hasattr(h, s), f"libf5e_b200.so does not export {s}"
    h.f5e_version.restype = ctypes.c_int
    assert h.f5e_version() >= 100


def test_no_cpu_fallback():
    import pytest
    from f5e_tts_b200.frontend import MelSpec
    from f5e_tts_b200.model import DiT
    from f5e_tts_b200.vocos import Vocos
    with pytest.raises(RuntimeError):
        MelSpec()(torch.zeros(1, 4096))
    with pytest.raises(RuntimeError):
        Vocos().decode(torch.zeros(1, 100, 10))
    dit = DiT(dim=256, depth=1, heads=4, text_dim=64, conv_layers=1, text_num_embeds=10)
    with pytest.raises(RuntimeError):
        dit.sample(torch.zeros(1, 8, 100), torch.zeros(1, 8, 100), torch.zeros(1, 3, dtype=torch.long), None, torch.tensor(0.5),
                   False, False, False)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "f5e_tts_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "f5e_oracle" not in src, f


def test_state_dict_keys_match_reference_layout():
    """a reference checkpoint must load unchanged."""
    from f5e_tts_b200.model import DiT
    from oracle import f5e_oracle as O
    cfg = O.DiTConfig(depth=2, use_ppg=True, use_codebook=True)
    ppg = dict(use_ppg=True, ppg_dim=256, use_transformer=False)
    cb = dict(use_codebook=True, num_vars=100, temp_start=2.0, temp_stop=0.5, temp_decay=0.999, groups=2, combine_groups=False,
              weight_proj_depth=1, weight_proj_factor=1)
    dit = DiT(dim=1024, depth=2, heads=16, ff_mult=2, text_dim=512, conv_layers=4, text_num_embeds=2545, ppg_config=ppg, cb_config=cb)
    sd = O.make_state_dict(cfg, 1)
    missing, unexpected = dit.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("num_batches_tracked" in k or "inv_freq" in k for k in missing), missing
    for k, v in sd.items():
        assert dit.state_dict()[k].shape == v.shape, k


def test_state_dict_keys_transformer_ppg_embedder():
    """PPGEmbedding(use_transformer=True) (dit.py:105-119): same keys as the reference's nn.TransformerEncoder; the oracle's
    key layout is pinned by tests/golden/dit_eval_small_ppg_transformer.pt (loaded into the reference with strict keys)."""
    from f5e_tts_b200.model import DiT
    from oracle import f5e_oracle as O
    cfg = O.DiTConfig(dim=512, depth=1, heads=8, text_dim=256, conv_layers=1, text_num_embeds=50, use_ppg=True, ppg_dim=256,
                      ppg_transformer=True, ppg_nhead=4, ppg_ff=512, ppg_layers=2)
    ppg = dict(use_ppg=True, ppg_dim=256, use_transformer=True,
               transformer_config=dict(nhead=4, dim_feedforward=512, dropout=0.1, num_layers=2))
    dit = DiT(dim=512, depth=1, heads=8, ff_mult=2, text_dim=256, conv_layers=1, text_num_embeds=50, ppg_config=ppg)
    sd = O.make_state_dict(cfg, 1)
    missing, unexpected = dit.load_state_dict(sd, strict=False)
    assert not unexpected and all("inv_freq" in k for k in missing), (missing, unexpected)
    with pytest.raises(NotImplementedError):  # head width other than 64 is not supported by the attention kernel
        DiT(dim=512, depth=1, heads=8, ppg_config=dict(use_ppg=True, ppg_dim=256, use_transformer=True,
                                                       transformer_config=dict(nhead=8, dim_feedforward=512, dropout=0.1, num_layers=1)))


def test_vocos_state_dict_layout():
    from f5e_tts_b200.vocos import Vocos
    from oracle import f5e_oracle as O
    sd = O.make_vocos_state_dict(1)
    sd["feature_extractor.mel_spec.window"] = torch.hann_window(1024)  # present in the published checkpoint, ignored
    v = Vocos()
    v.load_state_dict(sd)
    assert set(v.state_dict()) == {k for k in sd if not k.startswith("feature_extractor.")}


def test_shard_plan_is_balanced_and_complete():
    import numpy as np
    from f5e_tts_b200 import dist as fd
    rng = np.random.default_rng(20261104)
    secs = rng.uniform(3, 15, 1088)  # BASELINE configs[3]
    dur = [int(s * 24000 / 256) for s in secs]
    for world in (1, 2, 4, 8):
        shards = fd.plan_shards(dur, world)
        seen = sorted(i for s in shards for b in s for i in b.indices)
        assert seen == list(range(1088))
        for s in shards:
            for b in s:
                assert b.n_pad == max(dur[i] for i in b.indices) and len(b.indices) * b.n_pad <= 16 * 1031
        assert fd.imbalance(shards) <= 1.03, (world, fd.imbalance(shards))
    assert fd.plan_shards(dur, 4) == fd.plan_shards(dur, 4)  # deterministic: every rank derives the same plan


def test_shard_plan_edge_cases():
    from f5e_tts_b200 import dist as fd
    assert fd.plan_shards([], 2) == [[], []]
    s = fd.plan_shards([100], 4)
    assert sum(len(x) for x in s) == 1
    s = fd.plan_shards([4096] * 3, 2, frame_budget=4096)
    assert sorted(len(x) for x in s) == [1, 2]


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from f5e_tts_b200 import dist as fd
dist.init_process_group("gloo")
r = dist.get_rank()
t = torch.arange((r + 2) * 3, dtype=torch.float32).view(1, -1).repeat(2, 1) + 100 * r
g = fd.gather_padded(t, dst=0)
lens = torch.tensor([2 + r, 1])
pay = torch.arange(int(lens.sum()) * 4, dtype=torch.float32).view(-1, 4) + 1000 * r
al, ap = fd.gather_varlen(lens, pay, dst=0)
# split="dynamic": the ranks draw from one shared counter - together every batch exactly once, dearest first
from f5e_tts_b200 import evaluate as ev
prompts = [(None, None, None, None, [100 + 37 * (i % 5)] * (1 + i % 3), None) for i in range(23)]
mine = list(ev.split_batches(prompts, r, 2, "dynamic"))
got = [None, None]
dist.all_gather_object(got, mine)
assert sorted(got[0] + got[1]) == list(range(23)), got
cost = [sum(fd.utterance_cost(d) for d in p[4]) for p in prompts]
assert all(cost[a] >= cost[b] for a, b in zip(mine, mine[1:])), mine
if r == 0:
    assert g.shape == (4, 9), g.shape
    assert torch.equal(g[:2, :6], torch.arange(6.).repeat(2, 1)) and g[:2, 6:].abs().sum() == 0
    assert torch.equal(g[2:], torch.arange(9.).repeat(2, 1) + 100)
    assert al.tolist() == [2, 1, 3, 1] and ap.shape == (7, 4)
    assert torch.equal(ap[3:], torch.arange(16.).view(4, 4) + 1000)
    print("GLOO_OK")
else:
    assert g is None and al is None
dist.destroy_process_group()
"""


def test_gather_world_size_2_gloo(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(_GLOO_WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29613", str(w), ROOT], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0 and "GLOO_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def _header_struct_fields(name):
    """(field name, is_pointer) of a `typedef struct <name> {...}` in include/f5e_b200.h, in declaration order."""
    import re

    from f5e_tts_b200 import _lib

    src = re.sub(r"/\*.*?\*/", "", _lib.header_path().read_text(), flags=re.S)
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), src, flags=re.S).group(1)
    out = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        ptr_type = "*" in decl.split(",")[0]
        names = re.findall(r"[*\s,]([A-Za-z_][A-Za-z0-9_]*)\s*(?=,|$)", decl)
        out += [(n, ptr_type) for n in names]
    return out


def test_epilogue_struct_matches_header_and_docs():
    """The ctypes mirror, the header and the stub printed in INTEGRATION.md declare the same fields in the same order
    (a stale stub mis-lays the struct silently)."""
    import ctypes as C
    import re
    from pathlib import Path

    from f5e_tts_b200 import _lib

    hdr = _header_struct_fields("f5e_epilogue")
    mirror = [(n, t is C.c_void_p) for n, t in _lib.Epilogue._fields_]
    assert mirror == hdr
    doc = (Path(_lib.header_path()).parent.parent / "INTEGRATION.md").read_text()
    stub = doc[doc.index("class f5e_epilogue(C.Structure):"):]
    stub = stub[:stub.index("]\n") + 1]
    doc_fields = [(n, t == "c_void_p") for n, t in re.findall(r'\("([a-z0-9_]+)", C\.(c_void_p|c_int32)\)', stub)]
    assert doc_fields == hdr
    g = _header_struct_fields("f5e_guidance")
    assert [n for n, _ in g] == [n for n, _ in _lib.Guidance._fields_]


def test_library_stamp_matches_sources():
    """The loaded library carries the hash of the sources / flags it was built from; `_lib.load()` refuses a stale one."""
    import ctypes as C

    from f5e_tts_b200 import _lib, build

    lib = _lib.load()
    lib.f5e_build_stamp.restype = C.c_char_p
    assert lib.f5e_build_stamp().decode() == build.source_stamp()


def test_row_packing_index_plumbing():
    """make_packing (include/f5e_b200.h: f5e_dit_packing): integer plumbing in torch, device independent.  Dropping form: the
    valid rows of every sample back to back.  Split form (the exact default): the same prefix, then the padded rows, i.e. a
    permutation of all samples * n_pad rows; every guidance branch contributes the same number of valid and of padded rows."""
    import torch
    from f5e_tts_b200.model import make_packing
    n_pad, klen1 = 9, torch.tensor([9, 4, 7])
    klen = klen1.repeat(2)                                   # two guidance branches
    S = klen.numel()
    drop, keep_d = make_packing(klen, n_pad)
    split, keep_s = make_packing(klen, n_pad, keep_padded=True)
    valid = int(klen.sum())
    assert drop.rows_packed == split.rows_packed == valid and drop.rows_total == 0 and split.rows_total == S * n_pad
    row_map_d, row_info_d, row_start_d = keep_d
    row_map_s, row_info_s, row_start_s = keep_s
    assert torch.equal(row_map_s[:valid], row_map_d) and torch.equal(row_info_s[:valid], row_info_d)
    assert torch.equal(row_start_s, row_start_d) and row_start_d.tolist() == [0, 9, 13, 20, 29, 33]
    assert sorted(row_map_s.tolist()) == list(range(S * n_pad))                      # a permutation of all rows
    sample, frame = row_info_s >> 16, row_info_s & 0xFFFF
    assert torch.equal(sample * n_pad + frame, row_map_s)                            # row_info agrees with row_map
    assert bool((frame[:valid] < klen[sample[:valid].long()]).all()) and bool((frame[valid:] >= klen[sample[valid:].long()]).all())
    # f5e_ode_solve addresses the STATE rows with the first 1 / branches of each part: those must be branch 0's rows
    half_v, half_p = valid // 2, (S * n_pad - valid) // 2
    assert bool((sample[:half_v] < S // 2).all()) and bool((sample[valid: valid + half_p] < S // 2).all())
    assert bool((sample[half_v:valid] >= S // 2).all()) and bool((sample[valid + half_p:] >= S // 2).all())
