#!/bin/bash
# Developer tool (run under gpurun): the counters that DO track tcgen05 on sm_100 - math ops executed in the tensor path
# (sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32, counts UTCHMMA work), the HMMA sub-pipe's active cycles and instruction
# count, the XU (MUFU) and TMEM pipe instruction counts, DRAM bytes - for every hot kernel of tools/microbench.py.
#   bash tools/ncu_tensor.sh gpurun_out/r02_tensor.csv
set -e
OUT=${1:-gpurun_out/r02_tensor.csv}
export F5E_NO_BUILD=1
M="gpu__time_duration.sum,sm__cycles_elapsed.max,sm__cycles_elapsed.avg,sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32.sum,sm__inst_executed_pipe_tensor_subpipe_hmma.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg,sm__pipe_tensor_cycles_active.avg,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_tmem.sum,sm__inst_executed_pipe_uniform.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__cycles_active.avg"
python tools/microbench.py --iters 2 --warm 1 > gpurun_out/micro_plain2.log 2>&1 &&
ncu --metrics $M --clock-control none -k regex:"gemm_bf16_kernel|attention_kernel|conv_pos_kernel|norm_mod_kernel|row_stats" -c 60 --csv --log-file $OUT python tools/microbench.py --iters 2 --warm 1 > gpurun_out/micro_ncu.log 2>&1
