#!/bin/bash
mkdir -p gpurun_out
nproc; free -g | head -2
timeout 900 python -m pytest tests/test_gpu_frontends.py -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -15
timeout 2400 python -m pytest tests/test_gpu_baseline_dims.py -q -s --timeout 900 -p no:cacheprovider --durations=8 2>&1 | grep -v "^$" | tail -60
