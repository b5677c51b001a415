#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_block.py -m gpu -q -x -k "window_attn or block" > gpurun_out/t_attn.log 2>&1; echo "attn+block tests rc=$?"; tail -8 gpurun_out/t_attn.log
python scripts/attn_trace.py c1 > gpurun_out/trace2_c1.txt 2>&1; cat gpurun_out/trace2_c1.txt
python scripts/attn_trace.py c2 > gpurun_out/trace2_c2.txt 2>&1; head -4 gpurun_out/trace2_c2.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_c1.json'))
    print({k:d[k] for k in ('value','ms_per_step')}, d['roofline']['achieved'], d['roofline']['avg_launch_ms'], d['e2e']['value'], d['clocks'])
except Exception as e: print('bench parse failed', e)
PY
tail -3 gpurun_out/bench_c1.err
