#!/bin/bash
# Run 65: the 7B workloads with the final kernels (C3: 121 x 1080p; C4: 37-frame 4K chunk)
mkdir -p gpurun_out
for w in c3 c4; do
  timeout 700 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$w.json'))
print('$w', {k:round(d[k],3) for k in ('value','ms_per_step')}, 'e2e', round(d['e2e']['value'],3), 'attn', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'share', round(d['roofline']['share_of_step'],3), d['clocks']['sm_mhz'])
PY
done
