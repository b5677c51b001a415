#!/bin/bash
# Run 23: modnorm software-pipelined prefetch A/B (isolated, L2 flushed) + kernel parity
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "modnorm or gate_residual" 2>&1 | tail -2
for s in c1 c2; do
  echo "--- prefetch on, $s";  timeout 120 python scripts/prof_kernels.py $s 2>&1 | tail -5
  echo "--- prefetch off, $s"; SVR2_MODNORM_PREFETCH=0 timeout 120 python scripts/prof_kernels.py $s 2>&1 | tail -5
done
