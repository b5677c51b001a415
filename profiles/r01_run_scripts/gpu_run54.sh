#!/bin/bash
# Run 54: per-instruction warp-state sampling (ncu SourceCounters) of the two shipped attention kernels at C1
mkdir -p gpurun_out
for impl in v2 v6; do
SVR2_ATTN_IMPL=$impl timeout 120 python scripts/prof_attn.py c1 || exit 1
SVR2_ATTN_IMPL=$impl timeout 400 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --import-source on --clock-control none -k regex:window_attn_tc -s 2 -c 1 -f -o gpurun_out/attn_src_$impl python scripts/prof_attn.py c1 > gpurun_out/ncu_attn_src_$impl.log 2>&1; echo "ncu $impl rc=$?"; tail -1 gpurun_out/ncu_attn_src_$impl.log
done
ls -la gpurun_out/*.ncu-rep
