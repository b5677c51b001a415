#!/bin/bash
# Run 72: SwiGLU in-projections as one stacked-weight GEMM (SVR2_SWIGLU_STACKED=1) against two GEMMs: parity, then in-step A/B at C1
mkdir -p gpurun_out
SVR2_SWIGLU_STACKED=1 timeout 300 python -m pytest tests/test_gpu_block.py tests/test_gpu_stack.py -m gpu -q -x -k "3b" 2>&1 | tail -2
for f in 0 1 0 1; do
  SVR2_SWIGLU_STACKED=$f timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sw$f.json 2> gpurun_out/bench_sw$f.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_sw$f.json'))
print('stacked=$f', {k:round(d[k],3) for k in ('value','ms_per_step')}, d['clocks']['sm_mhz'], d['gpu_launches'])
PY
done
