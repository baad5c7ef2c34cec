"""Per-kernel counts of the Blackwell-only SASS mnemonics in the shipped library (B200_PROFILING.md "What proves a
Blackwell-native kernel"): UTC*MMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / .st), UTMALDG / UTMASTG / UTMAREDG / UBLKCP
(TMA), UTCBAR (tcgen05.commit), plus the legacy tensor path (HMMA) that must stay at zero.

    python tools/sass_summary.py > profiles/sass_summary.md         # needs cuobjdump; no GPU
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "f5e_tts_b200", "libf5e_b200.so")
PAT = collections.OrderedDict([
    ("UTC*MMA", re.compile(r"\bUTC[A-Z]*MMA(\.[A-Z0-9_]+)*")), ("UTCBAR", re.compile(r"\bUTCBAR")),
    ("LDTM", re.compile(r"\bLDTM")), ("STTM", re.compile(r"\bSTTM")),
    ("UTMALDG", re.compile(r"\bUTMALDG")), ("UTMASTG", re.compile(r"\bUTMASTG")), ("UTMAREDG", re.compile(r"\bUTMAREDG")),
    ("UBLKCP", re.compile(r"\bUBLKCP")), ("MUFU.EX2", re.compile(r"\bMUFU\.EX2")), ("HMMA (legacy)", re.compile(r"\bHMMA")),
    ("BRA.U.ANY (serialised issue)", re.compile(r"\bBRA\.U\.ANY")),
])


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
        return dict(zip(names, out))
    except OSError:
        return {n: n for n in names}


def short(name):
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name.replace("f5e::", "").replace("(anonymous namespace)::", "")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, variants, cur = collections.OrderedDict(), {}, None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            variants[cur] = collections.Counter()
            continue
        if cur is None or "/*" not in line:
            continue
        for key, pat in PAT.items():
            mm = pat.search(line)
            if mm:
                counts[cur][key] += 1
                if key == "UTC*MMA":
                    variants[cur][mm.group(0)] += 1
    dm = demangle(list(counts))
    arch = re.findall(r"arch = (sm_\w+)", sass)
    print("# SASS evidence: Blackwell-native instructions per kernel of the shipped `libf5e_b200.so`\n")
    print(f"`cuobjdump -sass f5e_tts_b200/libf5e_b200.so` ({', '.join(sorted(set(arch)))}), counted by `tools/sass_summary.py`.  "
          "`UTC*MMA` = `tcgen05.mma`, `LDTM`/`STTM` = `tcgen05.ld`/`.st`, `UTMALDG`/`UTMASTG`/`UTMAREDG` = TMA load / store / "
          "reduce-add, `UTCBAR` = `tcgen05.commit`.  `HMMA` (the `mma.sync` path) must be absent.  `BRA.U.ANY` marks a tensor / TMA "
          "instruction wrapped in a per-thread serialisation loop (what `if (lane == 0)` issuers compiled to in round 1; "
          "`elect.sync` issuers compile without it).\n")
    keys = list(PAT)
    print("| kernel | " + " | ".join(keys) + " | tcgen05.mma forms |")
    print("|---|" + "---:|" * len(keys) + "---|")
    tot = collections.Counter()
    for fn, c in counts.items():
        if not any(c[k] for k in keys if k != "MUFU.EX2"):
            continue
        tot.update(c)
        forms = ", ".join(f"{k} x{v}" for k, v in sorted(variants[fn].items()))
        print(f"| `{short(dm.get(fn, fn))}` | " + " | ".join(str(c[k]) for k in keys) + f" | {forms} |")
    print("| **total** | " + " | ".join(f"**{tot[k]}**" for k in keys) + " | |")
    n_k = len(counts)
    print(f"\n{n_k} kernels in the library; the ones not listed use none of the above (bandwidth-bound CUDA-core kernels).")
    if tot["HMMA (legacy)"]:
        print("\n**WARNING: legacy HMMA instructions present.**")
        sys.exit(1)


if __name__ == "__main__":
    main()
