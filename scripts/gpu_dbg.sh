#!/bin/bash
mkdir -p gpurun_out
V=$PWD/comfyui-seedvr2_b200/csrc/build/variants
for cfg in "c2 swin" "c1 swin" "c1 plain" "c3 swin"; do
  SVR2_LIB=$V/lib_dbg.so timeout 200 python scripts/attn_stress.py $cfg 400 2>&1 | grep -E "svr2:|ok|differs" | head -5
done
timeout 200 python scripts/attn_stress.py c2 swin 400 2>&1 | grep -E "svr2:|ok|differs|rror" | head -5
timeout 120 python scripts/attn_ab.py turns 2>&1 | tail -1
SVR2_LIB=$V/lib_noturns.so timeout 120 python scripts/attn_ab.py noturns 2>&1 | tail -1
