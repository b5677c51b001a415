#!/bin/bash
# Run 44: ncu of the shipped C1 attention kernel (generation 8): hardware counters incl. DRAM bytes, then source counters
mkdir -p gpurun_out
timeout 120 python scripts/prof_attn.py c1 > gpurun_out/prof_plain.txt 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.txt; exit 1; }
cat gpurun_out/prof_plain.txt
timeout 300 ncu --section SpeedOfLight --section SchedulerStats --section WarpStateStats --section ComputeWorkloadAnalysis --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:window_attn_tc -s 2 -c 1 -f -o gpurun_out/attn_g8_a python scripts/prof_attn.py c1 > gpurun_out/ncu_attn_g8_a.log 2>&1; echo "ncu (hw counters) rc=$?"; tail -1 gpurun_out/ncu_attn_g8_a.log
ls -la gpurun_out/attn_g8_a.ncu-rep
