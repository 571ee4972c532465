#!/bin/bash
# round 2, run 5: bisect of the tensor-path GEMM with parts switched off (compile-time B200P_TC_DEBUG builds), R = 128 whole tiles
mkdir -p gpurun_out
for m in 0 1 2 4 8 15; do
  if [ $m = 0 ]; then unset B200P_LIB; else export B200P_LIB=$PWD/build/dbg$m/libb200parallm.so; fi
  echo "== debug mask $m (1 no X loads, 2 no MMA issue, 4 no epilogue, 8 no A stores)"
  B200P_TC_R=128 timeout 200 python scripts/qmm_layer_bench.py 8,64 llama3-8b gateup,down,head 2>&1 | grep -v "^model"
done 2>&1 | tee gpurun_out/r2_run5_bisect.txt
