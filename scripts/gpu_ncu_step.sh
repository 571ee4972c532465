#!/bin/bash
# per-launch device times of one decode step (B=64, T=256 and T=8192) + full ncu capture of the gate|up GEMM
mkdir -p gpurun_out
W="--workload llama3-8b-4bit-b64-t256 --sweep= --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 python bench.py $W > gpurun_out/plain_t256.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'qmv|qmm_tc|attn_|rmsnorm|rope_kv|dequant|argmax|advance_len|silu_mul' -c 1200 --csv \
  --log-file gpurun_out/launches_t256.csv python bench.py $W > gpurun_out/ncu_t256.log 2>&1
echo "ncu launches exit: $?"
python scripts/tc_one.py 64 28672 4096 2 > gpurun_out/tc_one_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/prof_qmm_gateup_v3 python scripts/tc_one.py 64 28672 4096 2 > gpurun_out/ncu_qmm.log 2>&1
echo "ncu qmm exit $?"; cat gpurun_out/tc_one_plain.log
