#!/bin/bash
mkdir -p gpurun_out
python scripts/attn_trace.py c1 > gpurun_out/trace2_c1.txt 2>&1; cat gpurun_out/trace2_c1.txt
python scripts/attn_trace.py c2 > gpurun_out/trace2_c2.txt 2>&1; cat gpurun_out/trace2_c2.txt | tail -5
