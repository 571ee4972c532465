#!/bin/bash
# round 2, run 17: whole-tile plans on tensor-parallel shard shapes (one GPU, shapes only)
mkdir -p gpurun_out
for m in llama3-8b-tp2 llama3-8b-tp4 llama3-8b-tp8; do
  echo "=== $m default plan"
  timeout 200 python scripts/qmm_layer_bench.py 64 $m gateup,head 2>&1 | grep -v "^model"
  for V in "B200P_TC_R=128" "B200P_TC_R=104" "B200P_TC_R=64" "B200P_TC_S=1" "B200P_TC_S=2" "B200P_TC_S=4"; do
    echo "--- $V"
    env $V timeout 200 python scripts/qmm_layer_bench.py 64 $m gateup,head 2>&1 | grep -v "^model"
  done
done 2>&1 | tee gpurun_out/r2_run17_tp_shard_plans.txt
