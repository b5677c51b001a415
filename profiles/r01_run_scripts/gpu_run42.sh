#!/bin/bash
# Run 42: in-step A/B of attention generation 4 (v2) vs generation 8 (v6, cta_group::2), same box
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -1
for w in c1 c3r c2; do
for impl in v2 v6 v2 v6; do
  SVR2_ATTN_IMPL=$impl timeout 600 python bench.py --steps 4 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_${w}_$impl.json 2> gpurun_out/bench_${w}_$impl.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_${w}_$impl.json'))
print('$w $impl', {k:round(d[k],3) for k in ('value','ms_per_step')}, 'attn', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_ms']*1e3,1), d['clocks']['sm_mhz'])
PY
done
done
