#!/bin/bash
# Run 52: share of the exponentials evaluated on the FMA pipe (SVR2_FMA8 of every 8 pairs; shipped: 3), isolated launches
V=comfyui-seedvr2_b200/csrc/build/variants
for rep in 1 2; do
for impl in v6 v2; do
SVR2_ATTN_IMPL=$impl timeout 120 python scripts/attn_ab.py "$impl fma 3/8 (shipped)"
for f in 2 4 5; do
SVR2_ATTN_IMPL=$impl SVR2_LIB=$V/lib_${impl}_fma$f.so timeout 120 python scripts/attn_ab.py "$impl fma $f/8"
done
done
done
