#!/bin/bash
# Run 63: generation 8, two more spill-free register splits against the shipped 48/64/200
V=comfyui-seedvr2_b200/csrc/build/variants
for rep in 1 2; do
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_ab.py "v6 48/64/200 (shipped)"
for n in 3264208 4072200; do SVR2_ATTN_IMPL=v6 SVR2_LIB=$V/lib_v6_r$n.so timeout 120 python scripts/attn_ab.py "v6 regs $n"; done
done
