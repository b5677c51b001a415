#!/bin/bash
# Run 49: generation 8, m / l hand-over through named barriers (default build) against the mbarrier hand-over (variant)
V=comfyui-seedvr2_b200/csrc/build/variants
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -1
SVR2_ATTN_IMPL=v6 timeout 300 python scripts/attn_stress.py 2>&1 | tail -2
for i in 1 2 3; do
SVR2_ATTN_IMPL=v6 SVR2_LIB=$V/lib_v6_mbar.so timeout 120 python scripts/attn_ab.py "v6 mbarrier hand-over"
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_ab.py "v6 named-barrier hand-over"
done
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py "v2"
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_v5_prof.py c2 2>&1 | grep -E "kernel|CTA 0 warpgroup 0|softmax warpgroup 0|MMA warp"
