#!/bin/bash
# Run 48: generation 8 with the prefetching softmax cursor; v2 is the same-box control (run 42: v2 163-165 us, v6 151.5 us at C1)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -1
for s in c2 c1; do
SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_v5_prof.py $s 2>&1 | grep -E "kernel|CTA 0 warpgroup 0|softmax warpgroup 0|MMA warp"
done
for w in c1 c2; do
for impl in v2 v6 v2 v6; do
  SVR2_ATTN_IMPL=$impl timeout 600 python bench.py --steps 4 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_${w}_$impl.json 2> gpurun_out/bench_${w}_$impl.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_${w}_$impl.json'))
print('$w $impl', {k:round(d[k],3) for k in ('value','ms_per_step')}, 'attn', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_ms']*1e3,1), d['clocks']['sm_mhz'])
PY
done
done
