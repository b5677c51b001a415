#!/bin/bash
mkdir -p gpurun_out
for s in c1 c2; do timeout 120 python scripts/attn_v5_prof.py $s 2>&1 | tail -4; done
