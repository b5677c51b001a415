#!/bin/bash
# Run 61: register split, more points (v2, v6, v7)
V=comfyui-seedvr2_b200/csrc/build/variants
for rep in 1 2; do
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py "v2 72/72/184 (shipped)"
for n in 5656200 4864200 6464192; do SVR2_ATTN_IMPL=v2 SVR2_LIB=$V/lib_v2_r$n.so timeout 120 python scripts/attn_ab.py "v2 regs $n"; done
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_ab.py "v6 72/72/184 (shipped)"
for n in 5656200 4864200; do SVR2_ATTN_IMPL=v6 SVR2_LIB=$V/lib_v6_r$n.so timeout 120 python scripts/attn_ab.py "v6 regs $n"; done
SVR2_ATTN_IMPL=v7 timeout 120 python scripts/attn_ab.py "v7 72/72/184"
SVR2_ATTN_IMPL=v7 SVR2_LIB=$V/lib_v7_r5656200.so timeout 120 python scripts/attn_ab.py "v7 regs 5656200"
done
