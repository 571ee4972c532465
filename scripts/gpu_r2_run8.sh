#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -k "qmm" -x -q --timeout 120 -p no:cacheprovider 2>&1 | tail -5
echo "== layer bench (default plan)"
timeout 300 python scripts/qmm_layer_bench.py 1,8,64 llama3-8b 2>&1 | tee gpurun_out/r2_run8_qmm_layer_bench.txt
