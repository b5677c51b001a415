#!/bin/bash
# Run 58: generation 8, softmax warpgroups walk items / own key tiles directly (default build) against the generic cursor (variant)
V=comfyui-seedvr2_b200/csrc/build/variants
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -2
for s in "c1 win" "c1 swin" "c2 swin"; do SVR2_ATTN_IMPL=v6 timeout 300 python scripts/attn_stress.py $s 200 2>&1 | tail -1; done
for rep in 1 2 3; do
SVR2_ATTN_IMPL=v6 SVR2_LIB=$V/lib_v6_flat.so timeout 120 python scripts/attn_ab.py "v6 generic cursor"
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_ab.py "v6 item / own-tile walk"
done
