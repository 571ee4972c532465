#!/bin/bash
# run 32: section timers inside the dequant warps (timing build), without PDL
mkdir -p gpurun_out
export B200P_LIB=$PWD/build/dbgt/libb200parallm.so
export B200P_PDL=0
for a in "8 28672 4096 3" "64 28672 4096 3" "64 128256 4096 1"; do
  python scripts/tc_timing.py $a small
done 2>&1 | tee gpurun_out/r2_run32_tc_timing_sections.txt
