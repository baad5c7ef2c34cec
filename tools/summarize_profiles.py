"""Turn the ncu artefacts brought back in gpurun_out/ into the small, tracked summaries under profiles/.
    python tools/summarize_profiles.py r01
Inputs : gpurun_out/<tag>_launches.csv (ncu --metrics gpu__time_duration.sum launch list of `bench.py`)
         gpurun_out/<tag>_full.ncu-rep (ncu --set full capture of the top kernels)
Outputs: profiles/<tag>_launches.md, profiles/<tag>_kernels.md, profiles/traffic.json (DRAM bytes per launch, read by bench.py)
"""
import csv
import io
import json
import os
import re
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"


def short(name):
    name = re.sub(r"^void\s+", "", name)
    m = re.match(r"(?:f5e::)?([A-Za-z0-9_]+)(<[^>(]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:60]


def launches():
    p = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
    if not os.path.exists(p):
        return
    lines = [l for l in open(p) if not l.startswith("==")]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    agg = defaultdict(lambda: [0.0, 0])
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
        k = short(r["Kernel Name"])
        agg[k][0] += v
        agg[k][1] += 1
    tot = sum(v[0] for v in agg.values())
    out = [f"# {tag}: kernel launch list of `bench.py` (ncu --metrics gpu__time_duration.sum --clock-control none)", "",
           "Cold-cache, serialised per-launch times: compare SHARES, not absolutes.", "",
           f"{sum(v[1] for v in agg.values())} launches captured, {tot / 1e3:.2f} ms of device time.", "",
           "| kernel | launches | total us | share | avg us |", "|---|---:|---:|---:|---:|"]
    for k, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        out.append(f"| `{k}` | {n} | {us:.1f} | {us / tot:.3f} | {us / n:.1f} |")
    open(os.path.join(ROOT, "profiles", f"{tag}_launches.md"), "w").write("\n".join(out) + "\n")
    print("\n".join(out[:20]))


KEYS = [("gpu__time_duration.sum", "duration us"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs"), ("launch__shared_mem_per_block_dynamic", "dyn smem KB"),
        ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (of peak, all dtypes)"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU pipe %"),
        ("dram__bytes_read.sum", "dram read MB"), ("dram__bytes_write.sum", "dram write MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %")]


def full():
    p = os.path.join(ROOT, "gpurun_out", f"{tag}_full.ncu-rep")
    if not os.path.exists(p):
        return
    raw = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units = rows[0], rows[1]
    what = "one DiT evaluation, BASELINE config 2 shape" if re.fullmatch(r"r\d+", tag) else "see the capture command in DESIGN.md"
    out = [f"# {tag}: `ncu --set full --clock-control none` of the top kernels ({what})", "",
           "Numbers taken under the profiler explain the kernels; they are never bench values.", ""]
    traffic = defaultdict(list)
    for v in rows[2:]:
        d = dict(zip(h, v))
        u = dict(zip(h, units))
        name = short(d.get("Kernel Name", "?"))
        out.append(f"## `{name}`  grid {d.get('launch__grid_size')} x {d.get('launch__block_size')}")
        for k, label in KEYS:
            if k in d and d[k] not in ("", None):
                out.append(f"- {label}: {d[k]} {u.get(k, '')}".rstrip())
        try:
            rd, wr = float(d["dram__bytes_read.sum"].replace(",", "")), float(d["dram__bytes_write.sum"].replace(",", ""))
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u.get("dram__bytes_read.sum", "Mbyte"), 1e6)
            traffic[name.split("<")[0] if re.fullmatch(r"r\d+", tag) else name].append((rd + wr) * scale)
        except (KeyError, ValueError):
            pass
        out.append("")
    open(os.path.join(ROOT, "profiles", f"{tag}_full.md"), "w").write("\n".join(out) + "\n")
    tj = {k: sum(v) / len(v) for k, v in traffic.items()}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp) and not re.fullmatch(r"r\d+", tag):  # side captures (e.g. r01ppg) add to the table, never replace it
        old = json.load(open(tp))
        old.update(tj)
        tj = old
    tj["_note"] = f"mean dram__bytes_read.sum + dram__bytes_write.sum per launch, bytes, from gpurun_out/{tag}_full.ncu-rep"
    json.dump(tj, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(json.dumps(tj, indent=1))


launches()
full()
