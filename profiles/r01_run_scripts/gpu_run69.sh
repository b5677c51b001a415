#!/bin/bash
# Run 69: generation 4 with Q K^T(j+1) issued in two halves (upper half under the softmax of step j) — variant build
V=comfyui-seedvr2_b200/csrc/build/variants/lib_v2_split.so
SVR2_LIB=$V SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_stress.py c1 win 30 2>&1 | tail -2
SVR2_LIB=$V timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -2
for s in "c1 swin" "c2 win" "c2 swin" "c3 swin"; do SVR2_LIB=$V SVR2_ATTN_IMPL=v2 timeout 300 python scripts/attn_stress.py $s 200 2>&1 | tail -1; done
for rep in 1 2; do
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py "v2 (shipped)"
SVR2_LIB=$V SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py "v2 split QK"
done
