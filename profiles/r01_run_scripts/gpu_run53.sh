#!/bin/bash
# Run 53: generation 9 (window_attn_v7.cu): CTA pairs, two query tiles per CTA, no m / l hand-over.  Correctness first (bounded
# by timeouts: every mbarrier wait in the kernel is bounded and traps), then isolated timing against v2 / v6.
SVR2_ATTN_IMPL=v7 timeout 120 python scripts/attn_stress.py c1 win 30 2>&1 | tail -3
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -5
for s in "c1 win" "c1 swin" "c2 win" "c2 swin" "c3 swin"; do SVR2_ATTN_IMPL=v7 timeout 300 python scripts/attn_stress.py $s 200 2>&1 | tail -2; done
for rep in 1 2; do
for impl in v2 v6 v7; do SVR2_ATTN_IMPL=$impl timeout 120 python scripts/attn_ab.py "$impl"; done
done
