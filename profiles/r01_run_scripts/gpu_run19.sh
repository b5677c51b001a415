#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/prof_kernels.py > gpurun_out/elem_plain.txt 2>&1 || { echo "plain failed"; tail -3 gpurun_out/elem_plain.txt; exit 1; }
tail -1 gpurun_out/elem_plain.txt
timeout 300 ncu --section SpeedOfLight --section SchedulerStats --section WarpStateStats --section MemoryWorkloadAnalysis --section Occupancy --section SourceCounters --section InstructionStats --import-source on --clock-control none -k regex:"qkv_pack_staged" -s 1 -c 1 -f -o gpurun_out/qkv_staged python scripts/prof_kernels.py > gpurun_out/ncu_qkv_staged.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_qkv_staged.log
