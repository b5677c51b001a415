#!/bin/bash
# Run 57 (8 GPUs): the driver's N=8 bench launch + the temporal-segment stitch at 8 ranks
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "own arm rc=$?"; tail -1 gpurun_out/bench_n8.json | cut -c1-300
NP=8 bash scripts/gpu_run51.sh
