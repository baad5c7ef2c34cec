"""GPU: programmatic dependent launch must not change a bit.  The library reads F5E_PDL once per process (0 = every launch
fully serialised, 1 = every Euler-loop kernel with the programmatic-stream-serialization attribute; unset = on below 6144
rows only), so each setting runs in its own interpreter: the same solve - large enough that the default would leave the
attribute off, stream and cached-graph path, ragged batch with the split row packing - and the md5 of the result.
A kernel that touched global memory before its griddepcontrol.wait would show up here as a difference."""
import hashlib
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r"""
import hashlib, json, sys, torch
sys.path.insert(0, sys.argv[1])
from f5e_tts_b200.model import CFM, DiT
dev = torch.device("cuda:0")
dit = DiT(dim=1024, depth=2, heads=16, ff_mult=2, mel_dim=100, text_num_embeds=256, text_dim=512, conv_layers=4)
g = torch.Generator().manual_seed(1)
with torch.no_grad():
    for p in dit.parameters():
        p.copy_(torch.randn(p.shape, generator=g) * (0.02 if p.dim() > 1 else 0.1))
model = CFM(dit).to(dev).to(torch.bfloat16)
B = 4
ref_len = torch.randint(40, 80, (B,), generator=g).to(dev)
dur = torch.randint(900, 1031, (B,), generator=g).to(dev)          # 2 * 4 * ~1000 rows: above the automatic threshold
cond = torch.randn(B, int(ref_len.max()), 100, generator=g).to(dev)
text = torch.randint(0, 200, (B, 40), generator=g).to(dev)
out = {}
for graph in (False, True):
    model.use_cuda_graph = graph
    for rep in range(2):                                             # the second graph solve is a cached replay
        y, traj = model.sample(cond, text, None, dur, lens=ref_len, seed=3, steps=3, cfg_strength=2.0, sway_sampling_coef=-1.0)
        torch.cuda.synchronize()
        out["graph%d_rep%d" % (graph, rep)] = hashlib.md5(traj.float().cpu().numpy().tobytes()).hexdigest()
print("RESULT " + json.dumps(out))
"""


def _run(pdl):
    env = dict(os.environ, F5E_PDL=pdl)
    r = subprocess.run([sys.executable, "-c", _WORKER, ROOT], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def test_pdl_on_equals_off_bit_for_bit(cuda_dev):
    off, on = _run("0"), _run("1")
    assert len(set(off.values())) == 1, off          # stream == graph == replay
    assert on == off, (on, off)
