#!/bin/bash
# A/B timing of attention-kernel variants built by scripts/build_variant.sh
mkdir -p gpurun_out
V=comfyui-seedvr2_b200/csrc/build/variants
( unset SVR2_LIB; timeout 120 python scripts/attn_ab.py default 2>&1 | tail -1 )
for f in $V/lib_*.so; do
  n=$(basename $f .so); n=${n#lib_}
  SVR2_LIB=$PWD/$f timeout 120 python scripts/attn_ab.py $n 2>&1 | tail -1
done
( unset SVR2_LIB; timeout 120 python scripts/attn_ab.py default-again 2>&1 | tail -1 )
