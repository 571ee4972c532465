#!/bin/bash
# run 33: mma.sync GEMV with 1 / 2 / 3 steps of weights in flight per lane
mkdir -p gpurun_out
for D in 1 2 3; do
  echo "=== B200P_GEMV_DEPTH=$D"
  B200P_GEMV_DEPTH=$D timeout 200 python scripts/qmm_sweep.py 1,2,4,8 2>&1 | grep -v "^N K" | awk -F'|' '{print $1 "|" $3 "|" $4}' | tee -a gpurun_out/r33_gemv_depth_$D.txt
done
