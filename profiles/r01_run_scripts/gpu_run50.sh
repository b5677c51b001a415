#!/bin/bash
# Run 50: the one-step SR call kernels (sr_input, v_lerp endpoint) + full GPU suite
timeout 600 python -m pytest tests/test_gpu_sampler.py -m gpu -q -x 2>&1 | tail -15
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
