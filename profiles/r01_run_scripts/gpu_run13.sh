#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -3 gpurun_out/bench_c1.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_c1.json'))
    print({k:d[k] for k in ('value','ms_per_step','gpu_launches','ms_per_step_kernel_by_kernel')}, d['roofline']['achieved'], d['e2e']['value'], d['clocks'])
    for k,v in d['kernels'].items(): print(k, v)
except Exception as e: print('parse failed', e)
PY
