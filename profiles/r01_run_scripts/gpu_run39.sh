#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "(window_attn_vs_fp32 and v6) or generations_bit_identical" 2>&1 | tail -1
for r in 1 2; do
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py v2 2>&1 | tail -1
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_ab.py v6-ahead3 2>&1 | tail -1
SVR2_ATTN_IMPL=v6 SVR2_LIB=comfyui-seedvr2_b200/csrc/build/variants/lib_v6_ahead2.so timeout 120 python scripts/attn_ab.py v6-ahead2 2>&1 | tail -1
done
