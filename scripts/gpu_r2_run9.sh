#!/bin/bash
mkdir -p gpurun_out
export B200P_LIB=$PWD/build/dbgt/libb200parallm.so
for a in "8 28672 4096 3" "64 28672 4096 3" "8 128256 4096 1" "64 128256 4096 1" "64 4096 14336 2"; do
  python scripts/tc_timing.py $a small
done 2>&1 | tee gpurun_out/r2_run9_tc_timing.txt
