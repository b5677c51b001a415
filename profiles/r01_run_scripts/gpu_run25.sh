#!/bin/bash
# Run 25: ncu --set full captures of every kernel of the shipped library at C1 sizes (one launch each), after plain runs exited 0
mkdir -p gpurun_out
timeout 120 python scripts/prof_kernels.py c1 > gpurun_out/elem_plain.txt 2>&1 || { echo "plain elem failed"; tail -3 gpurun_out/elem_plain.txt; exit 1; }
timeout 120 python scripts/prof_attn.py c1 > gpurun_out/attn_plain.txt 2>&1 || { echo "plain attn failed"; tail -3 gpurun_out/attn_plain.txt; exit 1; }
cat gpurun_out/elem_plain.txt gpurun_out/attn_plain.txt
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"qkv_pack_kernel|modnorm_kernel|gate_residual_kernel|swiglu_kernel" -s 5 -c 5 -f -o gpurun_out/r25_elem python scripts/prof_kernels.py c1 > gpurun_out/ncu_r25_elem.log 2>&1; echo "ncu elem rc=$?"; tail -1 gpurun_out/ncu_r25_elem.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:window_attn_tc -s 2 -c 1 -f -o gpurun_out/r25_attn python scripts/prof_attn.py c1 > gpurun_out/ncu_r25_attn.log 2>&1; echo "ncu attn rc=$?"; tail -1 gpurun_out/ncu_r25_attn.log
ls -la gpurun_out/*.ncu-rep
