"""profiles/r02_kernels.md from the ncu counter pass of tools/ncu_tensor.sh (the counters that track tcgen05 on sm_100).

    python tools/summarize_tensor.py gpurun_out/r02_tensor.csv gpurun_out/micro_plain.log > profiles/r02_kernels.md
"""
import collections
import csv
import json
import re
import sys

LABELS = ["gemm qkv (+bias)", "gemm qkv + RoPE epilogue", "gemm qkv plain", "gemm out (+gate, residual reduce-add)",
          "gemm out, residual READ in the epilogue (scalar loads: the path the fused statistics replaced by a TMA-staged tile)",
          "gemm ff1 (+tanh-GELU)", "gemm ff1 plain", "gemm ff2 (+gate, residual reduce-add)",
          "gemm ff2, residual READ in the epilogue", "gemm in_proj (K = 712)",
          "gemm proj_out (N = 100, fp32 out)", "attention", "conv_pos", "norm_mod", "row_stats"]


def main():
    rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
    by = collections.OrderedDict()
    for r in rows:
        by.setdefault((int(r["ID"]), r["Kernel Name"]), {})[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    events = {}
    if len(sys.argv) > 2:
        for l in open(sys.argv[2]):
            if l.startswith("{"):
                d = json.loads(l)
                events[d["kernel"]] = d
    groups = []
    for (i, name), m in by.items():  # the microbench launches every kernel 3 times (1 warm-up + 2 timed) in a fixed order
        short = re.sub(r"\(.*", "", name).replace("void ", "").replace("f5e::", "")
        if not groups or groups[-1][0] != short or len(groups[-1][1]) == 3:
            groups.append((short, []))
        groups[-1][1].append(m)
    print("# r02: ncu counters that track tcgen05 (UTCHMMA) on sm_100, per hot kernel\n")
    print("`bash tools/ncu_tensor.sh` (B200, `--clock-control none`, config-2 shapes: M = 32 960 rows = 32 samples x 1030 frames).\n"
          "What the counters mean on this chip (checked against the launch arithmetic below):\n\n"
          "* `sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32.sum` stays **0** for every tcgen05 kernel: it counts the legacy HMMA path only.\n"
          "* `sm__inst_executed_pipe_tensor_subpipe_hmma.sum` counts **UTCHMMA instructions exactly** (QKV GEMM: 129 x 12 tiles x 64 = 99 072).\n"
          "* `sm__pipe_tensor_subpipe_hmma_cycles_active.avg / sm__cycles_elapsed.avg` is summed over the SM's four sub-pipes: **divide by 4** "
          "for the fraction of cycles the tensor pipe is busy at its nominal rate (8192 dense bf16 FLOP / clk / SM).  This is the number "
          "that agrees with CUDA-event TFLOP/s: QKV GEMM 99 072 UMMAs x 128 cycles / 74 pairs = 171 k busy cycles of 219 k = 0.78.  "
          "(`profiles/r01_kernels.md` quoted `sm__pipe_tensor_cycles_active...pct_of_peak_sustained_active`, which is normalised "
          "differently and read 10-30 % for the same kernels - that column should not have been labelled \"of peak\".)\n"
          "* XU utilisation = `sm__inst_executed_pipe_xu.sum` x 32 lanes / (16 lanes/clk/SM x 148 SMs x cycles).\n\n"
          "ncu serialises launches and the clock it ran at differs from a timed run (column `GHz` = cycles / duration), so compare "
          "FRACTIONS; the `events` column is the same kernel timed with CUDA events in the plain run just before.\n")
    print("| kernel | launch | us (ncu) | GHz | UMMA instr | tensor pipe busy | XU busy | DRAM MB | events: us, TFLOP/s or GB/s |")
    print("|---|---|---:|---:|---:|---:|---:|---:|---|")
    ev_keys = ["gemm_qkv", "gemm_qkv_rope_compute", "gemm_qkv_plain", "gemm_out", "gemm_out_residual_read", "gemm_ff1", "gemm_ff1_plain",
               "gemm_ff2", "gemm_ff2_residual_read", "gemm_in_proj",
               "gemm_proj_out", "attention", "conv_pos", "norm_mod", "row_stats"]
    for gi, (short, ms) in enumerate(groups):
        m = ms[-1]
        cyc, dur = m["sm__cycles_elapsed.avg"], m["gpu__time_duration.sum"] / 1e3
        tens = m["sm__pipe_tensor_subpipe_hmma_cycles_active.avg"] / cyc / 4
        xu = m["sm__inst_executed_pipe_xu.sum"] * 32 / (16 * 148 * cyc)
        dram = (m["dram__bytes_read.sum"] + m["dram__bytes_write.sum"]) / 1e6
        e = events.get(ev_keys[gi] if gi < len(ev_keys) else "", {})
        es = f"{e.get('ms', 0) * 1e3:.1f} us, " + (f"{e['tflops']} TFLOP/s" if "tflops" in e else f"{e.get('gbs', '-')} GB/s") if e else "-"
        label = LABELS[gi] if gi < len(LABELS) else ""
        print(f"| `{short}` | {label} | {dur:.1f} | {cyc / dur / 1e3:.2f} | {m['sm__inst_executed_pipe_tensor_subpipe_hmma.sum']:.0f} | "
              f"{tens:.3f} | {xu:.3f} | {dram:.1f} | {es} |")


if __name__ == "__main__":
    main()
