#!/bin/bash
# Profiling builds of the library with parts of qmm_tc_kernel switched off (B200P_TC_DEBUG bit mask, see qmm_tc.cu):
#   build/dbg<mask>/libb200parallm.so, selected at run time with B200P_LIB=...   (scripts/tc_bisect.py)
set -e
cd "$(dirname "$0")/../mlx_parallm_b200/csrc"
for m in "$@"; do
  out=../../build/dbg$m
  mkdir -p $out
  if [ "$m" = "t" ]; then def="-DB200P_TC_TIMING=1"; else def="-DB200P_TC_DEBUG=$m"; fi     # "t": per-role wait timers
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
       $def -c qmm_tc.cu -o $out/qmm_tc.o
  objs=$(ls ../../build/obj/*.o | grep -v qmm_tc.o)
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libb200parallm.so $objs $out/qmm_tc.o -lcuda
done
