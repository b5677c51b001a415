#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" > gpurun_out/t_attn.log 2>&1; echo "attn tests rc=$?"; tail -25 gpurun_out/t_attn.log
timeout 300 python -m pytest tests/test_gpu_block.py -m gpu -q > gpurun_out/t_block.log 2>&1; echo "block tests rc=$?"; tail -5 gpurun_out/t_block.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_c1.json'))
    print({k:d[k] for k in ('value','ms_per_step')}, d['roofline']['achieved'], d['roofline']['avg_launch_ms'], d['e2e']['value'])
    print({k:v['ms_per_step'] for k,v in d['kernels'].items()})
except Exception as e: print('bench parse failed', e)
PY
tail -5 gpurun_out/bench_c1.err
timeout 300 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_c2.json'))
    print({k:d[k] for k in ('value','ms_per_step')}, d['roofline']['achieved'], d['roofline']['avg_launch_ms'])
except Exception as e: print('bench parse failed', e)
PY
