#!/bin/bash
# correctness + timing of the tensor-path GEMM alone (short, hang-guarded)
mkdir -p gpurun_out
timeout 240 python scripts/tc_check.py ${TC_ARGS:-} > gpurun_out/tc_check.txt 2>&1
echo "tc_check exit: $?" >> gpurun_out/tc_check.txt
cat gpurun_out/tc_check.txt
