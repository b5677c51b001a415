#!/bin/bash
# Run 34: boustrophedon item dealing in the shipped attention kernel: parity, then same-box A/B against plain round robin
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -2
for r in 1 2; do
  SVR2_LIB=comfyui-seedvr2_b200/csrc/build/variants/lib_deal0.so timeout 120 python scripts/attn_ab.py round-robin 2>&1 | tail -1
  timeout 120 python scripts/attn_ab.py boustrophedon 2>&1 | tail -1
done
