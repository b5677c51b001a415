#!/bin/bash
# experiment: generation 8 without any MMA issued (garbage results) — does the softmax step get cheaper?
for s in c2; do
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_v5_prof.py $s 2>&1 | grep -E "kernel|CTA 0|softmax warpgroup 0|MMA warp"
echo "--- no MMA"
SVR2_ATTN_IMPL=v6 SVR2_LIB=comfyui-seedvr2_b200/csrc/build/variants/lib_v6_nomma.so timeout 120 python scripts/attn_v5_prof.py $s 2>&1 | grep -E "kernel|CTA 0|softmax warpgroup 0|MMA warp"
done
