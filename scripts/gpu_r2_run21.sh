#!/bin/bash
# per-role wait timers of qmm_tc_kernel WITHOUT programmatic dependent launch (run 9 had it on: CTA 0 of launch i+1 started
# on a free SM while launch i was still running, so its X / MMA waits contained the previous launch)
mkdir -p gpurun_out
export B200P_LIB=$PWD/build/dbgt/libb200parallm.so
export B200P_PDL=0
for a in "8 28672 4096 3" "16 28672 4096 3" "32 28672 4096 3" "64 28672 4096 3" "64 128256 4096 1"; do
  python scripts/tc_timing.py $a small
done 2>&1 | tee gpurun_out/r2_run21_tc_timing_nopdl.txt
