#!/bin/bash
mkdir -p gpurun_out
python scripts/tc_one.py 64 28672 4096 3 > gpurun_out/tc_one_gateup.txt 2>&1
python scripts/tc_one.py 64 4096 4096 2 > gpurun_out/tc_one_o.txt 2>&1
python scripts/tc_check.py --nocheck > gpurun_out/tc_time.txt 2>&1
python scripts/tc_one.py 64 28672 4096 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/prof_qmm_v1 python scripts/tc_one.py 64 28672 4096 3 > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
cat gpurun_out/tc_one_gateup.txt gpurun_out/tc_one_o.txt gpurun_out/tc_time.txt
