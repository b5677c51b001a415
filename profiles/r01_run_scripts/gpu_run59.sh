#!/bin/bash
# Run 59: ncu of generation 4 at C2 (the kernel dispatched for long sequences): warp-state sampling + tensor-pipe / DRAM counters
mkdir -p gpurun_out
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/prof_attn.py c2 || exit 1
SVR2_ATTN_IMPL=v2 timeout 500 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --section SpeedOfLight --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_elapsed --import-source on --clock-control none -k regex:window_attn_tc -s 2 -c 1 -f -o gpurun_out/attn_src_v2_c2 python scripts/prof_attn.py c2 > gpurun_out/ncu_attn_src_v2_c2.log 2>&1; echo "ncu rc=$?"; tail -1 gpurun_out/ncu_attn_src_v2_c2.log
