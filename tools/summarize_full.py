"""profiles/r02_full.md + profiles/traffic.json from the `ncu --set full` captures of tools/ncu_full.sh.

    python tools/summarize_full.py            # reads gpurun_out/r02_full_{attn,conv,norm,gemm}.csv (raw pages exported on the GPU box)
"""
import csv
import io
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LABELS = {  # launch order of tools/microbench.py per group (1 warm-up + 2 timed launches each: the LAST of every triple is used)
    "gemm": ["qkv (+bias)", "qkv + RoPE epilogue", "qkv plain", "out (+gate, in-place reduce-add)", "out, residual read in the epilogue",
             "ff1 (+tanh-GELU)", "ff1 plain", "ff2 (+gate, in-place reduce-add)", "ff2, residual read in the epilogue",
             "in_proj (K = 712)", "proj_out (N = 100)"],
    "attn": ["attention 32 x 16 heads x 1030 frames"], "conv": ["conv_pos k = 31, 16 groups"], "norm": ["norm_mod", "row_stats"]}
KEYS = [("gpu__time_duration.sum", "us"), ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots %"), ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU pipe %"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed", "smem wavefronts %"),
        ("lts__t_sectors.avg.pct_of_peak_sustained_elapsed", "L2 sectors %"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("dram__bytes_read.sum", "DRAM read MB"), ("dram__bytes_write.sum", "DRAM write MB")]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}


def short(name):
    return re.sub(r"\(.*", "", name).replace("void ", "").replace("f5e::", "")


def main():
    out = ["# r02: `ncu --set full --clock-control none` of the hot kernels at the config-2 shapes (tools/ncu_full.sh)", "",
           "Numbers taken under the profiler explain the kernels; they are never bench values.  The tensor-pipe figures that track "
           "tcgen05 are in `r02_kernels.md`; this table is the memory side (DRAM bytes per launch feed `traffic.json` and the bench "
           "line's `traffic` fields) and the issue / XU / occupancy picture.", "",
           "| kernel | launch | " + " | ".join(k[1] for k in KEYS) + " |", "|---|---|" + "---:|" * len(KEYS)]
    traffic = {}
    for grp, labels in LABELS.items():
        p = os.path.join(ROOT, "gpurun_out", f"r02_full_{grp}.csv")
        if not os.path.exists(p):
            continue
        raw = open(p).read()
        rows = list(csv.reader(io.StringIO(raw)))
        h, units = rows[0], rows[1]
        recs = [dict(zip(h, r)) for r in rows[2:]]
        u = dict(zip(h, units))
        for i, lab in enumerate(labels):
            idx = 3 * i + 2
            if idx >= len(recs):
                break
            d = recs[idx]
            vals = []
            for k, _ in KEYS:
                try:
                    v = float(d[k].replace(",", "")) * SCALE.get(u.get(k, ""), 1.0)
                    if "bytes" in k:
                        v /= 1e6
                    vals.append(f"{v:.1f}")
                except (KeyError, ValueError):
                    vals.append("-")
            out.append(f"| `{short(d['Kernel Name'])}` | {lab} | " + " | ".join(vals) + " |")
            try:
                tb = sum(float(d[k].replace(",", "")) * SCALE.get(u.get(k, ""), 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            except (KeyError, ValueError):
                continue
            name = short(d["Kernel Name"]).split("<")[0]
            if grp == "gemm":
                if i in (0, 3, 5, 7):  # the four DiT-block GEMMs as the step runs them
                    traffic.setdefault("gemm_bf16_kernel", []).append(tb)
            else:
                traffic.setdefault(name, []).append(tb)
    open(os.path.join(ROOT, "profiles", "r02_full.md"), "w").write("\n".join(out) + "\n")
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    tj = json.load(open(tp)) if os.path.exists(tp) else {}
    for k, v in traffic.items():
        tj[k] = sum(v) / len(v)
    tj["_note"] = ("mean dram__bytes_read.sum + dram__bytes_write.sum per launch, bytes: DiT kernels from gpurun_out/r02_full_*.ncu-rep "
                   "(tools/ncu_full.sh; gemm = mean of qkv / out / ff1 / ff2), PPG kernels from round 1's r01ppg capture")
    json.dump(tj, open(tp, "w"), indent=1)
    print("\n".join(out[-16:]))
    print(json.dumps(tj, indent=1))


if __name__ == "__main__":
    main()
