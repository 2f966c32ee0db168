#!/bin/bash
# ncu counters of k_extend / k_shadow over one frame for experimental builds: tools/ncu_pair_fetch.sh <variant.so>...
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__data_pipe_lsu_wavefronts.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_set_accesses_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__warps_active.avg.pct_of_peak_sustained_active
for v in "$@"; do
    n=$(basename $v .so)
    PT_LIB_PATH=$PWD/$v timeout 300 ncu --metrics $M --section WarpStateStats --clock-control none -k regex:'k_extend|k_shadow' -c 10 --csv --page raw --log-file gpurun_out/ncu_pf_$n.csv python tools/profile_frame.py atrium 1 > gpurun_out/ncu_pf_$n.log 2>&1
    echo "$n rc=$?"
done
