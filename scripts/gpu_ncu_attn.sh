#!/bin/bash
# ncu captures of the attention kernel at C1: hardware-counter sections and source-level sampling, each after a plain run
mkdir -p gpurun_out
tag=${1:-g5}
timeout 120 python scripts/prof_attn.py c1 > gpurun_out/prof_plain.txt 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.txt; exit 1; }
cat gpurun_out/prof_plain.txt
timeout 150 ncu --section SpeedOfLight --section SchedulerStats --section WarpStateStats --section ComputeWorkloadAnalysis --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --section WorkloadDistribution --clock-control none -k regex:window_attn_tc -s 2 -c 1 -f -o gpurun_out/attn_${tag}_a python scripts/prof_attn.py c1 > gpurun_out/ncu_attn_${tag}_a.log 2>&1; echo "ncu (hw counters) rc=$?"; tail -1 gpurun_out/ncu_attn_${tag}_a.log
timeout 200 ncu --section SourceCounters --import-source on --clock-control none -k regex:window_attn_tc -s 2 -c 1 -f -o gpurun_out/attn_${tag}_b python scripts/prof_attn.py c1 > gpurun_out/ncu_attn_${tag}_b.log 2>&1; echo "ncu (source counters) rc=$?"; tail -1 gpurun_out/ncu_attn_${tag}_b.log
