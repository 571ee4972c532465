#!/bin/bash
# gpu_check + a per-launch ncu timing list of the weights-dominated decode step (B=64, T=256).
bash scripts/gpu_check.sh
W="--workload llama3-8b-4bit-b64-t256 --sweep= --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 python bench.py $W > gpurun_out/plain_t256.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'qmv|qmm_tc|attn_|rmsnorm|rope_kv|dequant|argmax|advance_len' -c 900 --csv \
  --log-file gpurun_out/launches_t256.csv python bench.py $W > gpurun_out/ncu_t256.log 2>&1
echo "ncu exit: $?"
