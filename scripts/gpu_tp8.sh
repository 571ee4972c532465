#!/bin/bash
# 8-GPU pass kept short (charged 8x): quick parity in both exchange modes, exchange microbenchmark, bench in both modes.
export TP_CHECK_QUICK=1
for mode in rows oneshot; do
  echo "== exchange mode $mode"
  B200P_TP_EXCHANGE=$mode WORKLOADS= bash scripts/gpu_tp.sh 8 2>&1 | grep -E "^\[tp|exit"
done
export SKIP_CHECK=1
XBENCH=1 WORKLOADS= bash scripts/gpu_tp.sh 8 2>&1 | grep -E "^tp|exit"
for mode in rows oneshot; do
  for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b64-t8192; do
    B200P_TP_EXCHANGE=$mode WORKLOADS=$W bash scripts/gpu_tp.sh 8 > /dev/null 2>&1
    cp gpurun_out/bench_tp8_$W.json gpurun_out/bench_tp8_${mode}_$W.json
    echo "$mode $W $(grep -o 'ms_per_step": [0-9.]*' gpurun_out/bench_tp8_$W.json | head -1)"
  done
done
