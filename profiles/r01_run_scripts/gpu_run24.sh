#!/bin/bash
# Run 24: 7B MLP with the cuBLASLt bias+GELU epilogue — parity (7B block + stack golden tests) and A/B at C3 per-rank size
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_block.py tests/test_gpu_stack.py -m gpu -q -x -s -k "7b" 2>&1 | grep -E "PSNR|block vid|passed|failed|Error" | head
for e in 0 1; do
  SVR2_GELU_EPILOGUE=$e timeout 600 python bench.py --steps 3 --warmup 3 --workload c3r --no-cpu-baseline > gpurun_out/bench_c3r_gelu$e.json 2> gpurun_out/bench_c3r_gelu$e.err; echo "bench c3r gelu_epilogue=$e rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_c3r_gelu$e.json'))
print({k:d[k] for k in ('value','ms_per_step')}, 'attn', round(d['roofline']['achieved'],1), d['roofline']['frac'], d['clocks']['sm_mhz'])
PY
done
