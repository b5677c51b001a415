#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/t_gpu.log
timeout 120 python scripts/attn_ab.py auto 2>&1 | tail -1
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py gen4 2>&1 | tail -1
SVR2_ATTN_IMPL=v4 timeout 120 python scripts/attn_ab.py gen6 2>&1 | tail -1
for cfg in "c2 swin" "c1 swin" "c1 plain"; do timeout 200 python scripts/attn_stress.py $cfg 300 2>&1 | grep -E "svr2:|ok|differs|rror" | head -3; done
