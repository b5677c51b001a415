#!/bin/bash
# Run 60: generation 4, register split control / epilogue / softmax (setmaxnreg): shipped 72/72/184 against variants
V=comfyui-seedvr2_b200/csrc/build/variants
for rep in 1 2; do
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py "v2 72/72/184 (shipped)"
for n in 5656200 4072200 4056208; do
SVR2_ATTN_IMPL=v2 SVR2_LIB=$V/lib_v2_r$n.so timeout 120 python scripts/attn_ab.py "v2 regs $n"
done
done
