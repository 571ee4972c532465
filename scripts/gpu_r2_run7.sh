#!/bin/bash
# round 2, run 7: ncu of gate|up M=64 (whole tiles R=128) with source-level sampling
mkdir -p gpurun_out
export B200P_TC_R=128
python scripts/tc_one.py 64 28672 4096 3 small > gpurun_out/tc_one_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/r2_run7_prof_qmm_gateup python scripts/tc_one.py 64 28672 4096 3 small > gpurun_out/ncu_qmm.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_qmm.log; cat gpurun_out/tc_one_plain.log
