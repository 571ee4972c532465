#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_lora.py tests/test_gpu_frontends.py tests/test_gpu_model.py -x -q --timeout 300 -p no:cacheprovider 2>&1 | tail -8
