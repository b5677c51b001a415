#!/bin/bash
# Run 32: in-step A/B of attention generation 4 vs 7 at C1 (same box), new full-size property cases
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "full_size" 2>&1 | tail -2
for impl in v2 v5 v2 v5; do
  SVR2_ATTN_IMPL=$impl timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1_$impl.json 2> gpurun_out/bench_c1_$impl.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_c1_$impl.json'))
print('$impl', {k:round(d[k],3) for k in ('value','ms_per_step')}, 'attn', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_ms']*1e3,1), d['clocks']['sm_mhz'])
PY
done
