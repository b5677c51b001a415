#!/bin/bash
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
timeout 120 python scripts/prof_kernels.py 2>&1 | tail -4
