#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn_vs_fp32 and v5" 2>&1 | tail -3
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "generations_bit_identical" 2>&1 | grep -E "AssertionError|passed|failed" | head
for impl in v2 v5; do SVR2_ATTN_IMPL=$impl timeout 120 python scripts/attn_ab.py $impl 2>&1 | tail -2; done
