#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "(window_attn_vs_fp32 and v6) or generations_bit_identical" 2>&1 | tail -2
for impl in v2 v6; do SVR2_ATTN_IMPL=$impl timeout 120 python scripts/attn_ab.py $impl 2>&1 | tail -1; done
for s in c1 c2; do SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_v5_prof.py $s 2>&1 | tail -6; done
