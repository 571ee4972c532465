#!/bin/bash
# ncu --set full of the tensor-path GEMM (gate|up shape, M=64) and of the decode-attention kernel
mkdir -p gpurun_out
python scripts/tc_one.py 64 28672 4096 3 > gpurun_out/tc_one_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/prof_qmm_gateup python scripts/tc_one.py 64 28672 4096 3 > gpurun_out/ncu_qmm.log 2>&1
echo "ncu qmm exit $?"
cat gpurun_out/tc_one_plain.log
