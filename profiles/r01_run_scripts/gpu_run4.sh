#!/bin/bash
mkdir -p gpurun_out
python scripts/attn_trace.py c1 > gpurun_out/trace_c1.txt 2>&1; echo "trace rc=$?"; cat gpurun_out/trace_c1.txt
python scripts/attn_trace.py c2 swin > gpurun_out/trace_c2.txt 2>&1; cat gpurun_out/trace_c2.txt
python scripts/prof_kernels.py > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"qkv_pack|modnorm" -s 8 -c 3 -o gpurun_out/elem_c1 python scripts/prof_kernels.py > gpurun_out/ncu_elem.log 2>&1; echo "ncu rc=$?"
cat gpurun_out/prof_plain.log
