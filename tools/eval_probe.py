"""Developer tool: the configs[3]-shaped eval set (bench.eval_set_leg) on one GPU for several batch formations.
    python tools/eval_probe.py 16:200 32:200 32:50 64:50      # <utterances of 1031 frames per batch>:<length buckets>
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from f5e_tts_b200.vocos import Vocos  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    model, _ = bench.build_model(dev)
    model.store_trajectory = False
    voc = Vocos(seed=4321).to(dev)
    for spec in sys.argv[1:]:
        b, k = spec.split(":")
        r = bench.eval_set_leg(model, voc, dev, 0, 1, 1088, batch_frames=int(b) * 1031, num_buckets=int(k))
        print(json.dumps(dict(spec=spec, audio_s_per_s=round(r["audio_s_per_s"], 1), solve_s=round(r["solve_s_max_over_ranks"], 2),
                              workload=r["workload"][60:140])), flush=True)


if __name__ == "__main__":
    main()
