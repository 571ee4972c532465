#!/bin/bash
# Round-1 run 5: tests + smoke + bench, then the ncu launch list of the headline command and full captures of the
# two dominant kernels (decode attention at T=8192, gate|up GEMM at M=64).
bash scripts/gpu_check.sh
W="--sweep= --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'qmv|qmm_tc|attn_|rmsnorm|rope_kv|dequant|argmax|advance_len|silu_mul|embed' -c 1500 --csv \
  --log-file gpurun_out/launches_headline.csv python bench.py $W > gpurun_out/ncu_headline.log 2>&1
echo "ncu launches exit: $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 100 -c 1 \
  -o gpurun_out/prof_attn_t8192 python bench.py $W > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn exit: $?"
python scripts/tc_one.py 64 28672 4096 2 > gpurun_out/tc_one_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/prof_qmm_gateup_v4 python scripts/tc_one.py 64 28672 4096 2 > gpurun_out/ncu_qmm.log 2>&1
echo "ncu qmm exit $?"; cat gpurun_out/tc_one_plain.log
