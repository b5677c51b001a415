#!/bin/bash
# Run 62: new register splits shipped (v2 56/56/200, v6 48/64/200): correctness, stress, in-step bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" 2>&1 | tail -2
for s in "c1 win" "c1 swin" "c2 swin" "c3 win"; do timeout 300 python scripts/attn_stress.py $s 200 2>&1 | tail -1; done
SVR2_ATTN_IMPL=v7 timeout 300 python scripts/attn_stress.py c2 swin 100 2>&1 | tail -1
for w in c1 c2; do
  timeout 600 python bench.py --steps 5 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$w.json'))
print('$w', {k:round(d[k],3) for k in ('value','ms_per_step')}, 'e2e', round(d['e2e']['value'],3), 'attn', d['roofline']['kernel'][-40:], round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_ms']*1e3,1), d['clocks']['sm_mhz'])
PY
done
