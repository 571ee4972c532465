#!/bin/bash
# round 2, run 14: the weights-bound point (B=64, T=256): ncu launch list of the step + full capture of decode attention
mkdir -p gpurun_out
W="--workload llama3-8b-4bit-b64-t256 --sweep= --steps 2 --warmup 3 --no-cpu-baseline --no-prefill"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'qmv|qmm_tc|qmm_sk|attn_|rmsnorm|rope_kv|dequant|argmax|advance_len|silu_mul|embed' -c 1500 --csv \
  --log-file gpurun_out/r2_run14_launches_b64_t256.csv python bench.py $W > gpurun_out/r2_run14_ncu.log 2>&1
echo "ncu launches exit: $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 100 -c 1 \
  -o gpurun_out/r2_run14_prof_attn_t256 python bench.py $W > gpurun_out/r2_run14_ncu_attn.log 2>&1; echo "ncu attn exit: $?"
