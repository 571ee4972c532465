#!/bin/bash
# Round-end evidence pass on one GPU: tests, smoke, the default bench command, its ncu launch list, full captures of the
# dominant kernels, quantiser throughput.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -4 > gpurun_out/final_pytest.txt; cat gpurun_out/final_pytest.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; wc -l gpurun_out/final_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "reference arm exit $?"; head -c 400 gpurun_out/final_bench_reference.json; echo
W="--sweep= --steps 2 --warmup 3 --no-cpu-baseline --no-prefill"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'qmv|qmm_tc|attn_|rmsnorm|rope_kv|dequant|argmax|advance_len|silu_mul|embed' -c 1500 --csv \
  --log-file gpurun_out/final_launches_headline.csv python bench.py $W > gpurun_out/final_ncu_headline.log 2>&1
echo "ncu launches exit: $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 100 -c 1 \
  -o gpurun_out/final_prof_attn_t8192 python bench.py $W > gpurun_out/final_ncu_attn.log 2>&1; echo "ncu attn exit: $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/final_prof_qmm_gateup python scripts/tc_one.py 64 28672 4096 2 > gpurun_out/final_ncu_qmm.log 2>&1; echo "ncu qmm exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qmv_mma -s 40 -c 1 -o gpurun_out/final_prof_qmv_mma_b1 python bench.py --workload llama3-8b-4bit-b1-t8192 $W > gpurun_out/final_ncu_qmv.log 2>&1; echo "ncu qmv_mma exit $?"
timeout 300 python scripts/quantize_bench.py 2>&1 | tail -6
