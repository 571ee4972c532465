#!/bin/bash
# Round-2 FINAL evidence pass on one GPU (final build): tests (timed), smoke, the default bench command and the reference arm as the driver runs
# them, the ncu launch list of the same command, full captures of the dominant kernels.
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider 2>&1 | tail -4 ) > gpurun_out/r2_final_pytest.txt 2>&1; cat gpurun_out/r2_final_pytest.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
( time timeout 900 python bench.py > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err ) 2>&1 | tail -3; echo "bench exit $?"; wc -l gpurun_out/r2_final_bench.json
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2_final_bench_reference.json 2> gpurun_out/r2_final_bench_reference.err ) 2>&1 | tail -3; head -c 600 gpurun_out/r2_final_bench_reference.json; echo
W="--sweep= --steps 2 --warmup 3 --no-cpu-baseline --no-prefill"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'qmv|qmm_tc|qmm_sk|attn_|rmsnorm|rope_kv|dequant|argmax|advance_len|silu_mul|embed' -c 1500 --csv \
  --log-file gpurun_out/r2_final_launches_headline.csv python bench.py $W > gpurun_out/r2_final_ncu_headline.log 2>&1
echo "ncu launches exit: $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 100 -c 1 \
  -o gpurun_out/r2_final_prof_attn_t8192 python bench.py $W > gpurun_out/r2_final_ncu_attn.log 2>&1; echo "ncu attn exit: $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/r2_final_prof_qmm_tc_gateup python scripts/tc_one.py 64 28672 4096 3 small > gpurun_out/r2_final_ncu_qmm.log 2>&1; echo "ncu qmm_tc exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qmm_sk -s 5 -c 1 -o gpurun_out/r2_final_prof_qmm_sk_down python scripts/tc_one.py 64 4096 14336 2 small > gpurun_out/r2_final_ncu_qmm_sk.log 2>&1; echo "ncu qmm_sk exit $?"
