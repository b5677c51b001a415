#!/bin/bash
# Run 67: ncu of qkv_pack at C1 (instruction mix, issue utilisation, stall sampling)
mkdir -p gpurun_out
timeout 120 python scripts/prof_kernels.py > gpurun_out/elem_plain.txt 2>&1 || { echo "plain failed"; tail -3 gpurun_out/elem_plain.txt; exit 1; }
grep -i "qkv_pack" gpurun_out/elem_plain.txt
timeout 300 ncu --section SpeedOfLight --section SchedulerStats --section WarpStateStats --section Occupancy --section SourceCounters --section InstructionStats --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_elapsed --import-source on --clock-control none -k regex:"qkv_pack_kernel" -s 2 -c 1 -f -o gpurun_out/qkv_pack_c1 python scripts/prof_kernels.py > gpurun_out/ncu_qkv.log 2>&1; echo "ncu rc=$?"; tail -1 gpurun_out/ncu_qkv.log
