#!/bin/bash
# Developer tool (run under gpurun): `ncu --set full` of the hot kernels at the config-2 shapes, one capture per microbench
# group (each preceded by the same command without ncu, as the profiling recipe asks), then the counter pass of
# tools/ncu_tensor.sh.  Reports land in gpurun_out/r02_full_<group>.ncu-rep; tools/summarize_full.py reads them.
export F5E_NO_BUILD=1
python tools/microbench.py > gpurun_out/micro_events.log 2>&1   # the CUDA-event numbers (20 launches each) beside the captures
for g in attn conv norm gemm; do
  c=3; [ $g = norm ] && c=6; [ $g = gemm ] && c=33
  python tools/microbench.py --only $g --iters 2 --warm 1 > gpurun_out/micro_plain_$g.log 2>&1 &&
  ncu --set full --clock-control none -k regex:"gemm_bf16_kernel|attention_kernel|conv_pos_kernel|norm_mod_kernel|row_stats_kernel" \
      -c $c -f -o gpurun_out/r02_full_$g python tools/microbench.py --only $g --iters 2 --warm 1 > gpurun_out/micro_ncu_$g.log 2>&1
  echo "$g rc=$?"
  # the raw page as CSV is what tools/summarize_full.py reads; the report itself is too big to bring back (64 MiB limit)
  ncu -i gpurun_out/r02_full_$g.ncu-rep --page raw --csv > gpurun_out/r02_full_$g.csv 2>/dev/null
  [ $g = attn ] || rm -f gpurun_out/r02_full_$g.ncu-rep
done
bash tools/ncu_tensor.sh gpurun_out/r02_tensor.csv; echo "tensor rc=$?"
