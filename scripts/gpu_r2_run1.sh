#!/bin/bash
# round 2, run 1: new stream-K fix-up (no atomics) + small-scale dequant: kernel tests, whole suite, GEMM layer bench,
# decode sweep, ncu of the gate|up GEMM with sources
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 600 python -m pytest tests/test_gpu_kernels.py -k "qmm" -x -q --timeout 120 -p no:cacheprovider 2>&1 | tail -15
echo "== full suite"
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -15
echo "== layer bench"
timeout 300 python scripts/qmm_layer_bench.py 1,8,64 llama3-8b 2>&1 | tee gpurun_out/r2_run1_qmm_layer_bench.txt
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b64-t8192 llama3-8b-4bit-b8-t8192 llama3-8b-4bit-b1-t8192; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r2_run1_bench_$W.json 2> gpurun_out/r2_run1_bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/r2_run1_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3), 'launches', d['gpu_launches']//16, d['clocks']); print({k: v['mean_us'] for k, v in d['kernels'].items()})"
done
echo "== ncu gate|up"
python scripts/tc_one.py 64 28672 4096 3 small > gpurun_out/tc_one_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:qmm_tc -s 5 -c 1 -o gpurun_out/r2_run1_prof_qmm_gateup python scripts/tc_one.py 64 28672 4096 3 small > gpurun_out/ncu_qmm.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_qmm.log; cat gpurun_out/tc_one_plain.log
