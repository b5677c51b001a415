#!/bin/bash
# Run 68 (4 GPUs): the driver's N=4 bench launch
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "own arm rc=$?"; tail -1 gpurun_out/bench_n4.json | cut -c1-260
