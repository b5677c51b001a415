#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -3 gpurun_out/bench_c1.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_c1.json'))
    print({k:d[k] for k in ('value','ms_per_step','gpu_launches','launch_mode','ms_per_step_kernel_by_kernel')}, d['roofline']['achieved'], d['e2e'], d['clocks'])
except Exception as e: print('parse failed', e)
PY
timeout 300 python -m pytest tests/test_gpu_block.py -m gpu -q -x 2>&1 | tail -2
