#!/bin/bash
# run 31: bisect of qmm_tc_kernel with TWO MMA issuers (compile-time B200P_TC_DEBUG builds; 1 no X loads, 2 no MMA issue,
# 4 no epilogue, 8 no dequant math / A stores); planned tile rows (R = 104 for gate|up, 112 for the head)
mkdir -p gpurun_out
for m in 0 2 4 8 6 10 12 14 15; do
  if [ $m = 0 ]; then unset B200P_LIB; else export B200P_LIB=$PWD/build/dbg$m/libb200parallm.so; fi
  echo "== debug mask $m"
  timeout 200 python scripts/qmm_layer_bench.py 8,64 llama3-8b gateup,head 2>&1 | grep -v "^model"
done 2>&1 | tee gpurun_out/r2_run31_qmm_bisect_2issuers.txt
