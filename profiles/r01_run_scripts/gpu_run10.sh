#!/bin/bash
mkdir -p gpurun_out
for v in 0 1; do
  if [ $v = 0 ]; then unset SVR2_LIB; else export SVR2_LIB=$PWD/comfyui-seedvr2_b200/csrc/build/variants/lib_exp$v.so; fi
  echo "=== variant $v"
  timeout 120 python scripts/attn_trace.py c1 2>&1 | grep -E "TFLOP|fine|item 3|rror"
done
