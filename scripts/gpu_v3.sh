#!/bin/bash
# v3 tensor-path GEMM: correctness (tc_check + pytest) and timing
mkdir -p gpurun_out
timeout 300 python scripts/tc_check.py > gpurun_out/tc_check.txt 2>&1; echo "tc_check exit $?" >> gpurun_out/tc_check.txt
tail -30 gpurun_out/tc_check.txt
timeout 600 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/pytest_gpu.txt
cat gpurun_out/pytest_gpu.txt
