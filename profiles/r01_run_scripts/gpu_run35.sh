#!/bin/bash
# Run 35: attention generation 8 (cta_group::2, SVR2_ATTN_IMPL=v6): parity on the small grids first, then A/B
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn_vs_fp32 and v6" 2>&1 | tail -6
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "generations_bit_identical" 2>&1 | grep -E "Error|passed|failed" | head -5
for impl in v2 v6; do SVR2_ATTN_IMPL=$impl timeout 120 python scripts/attn_ab.py $impl 2>&1 | tail -2; done
