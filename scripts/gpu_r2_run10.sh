#!/bin/bash
# round 2, run 10: full GPU suite + decode sweep with the whole-tile GEMM plan
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -8
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b64-t8192 llama3-8b-4bit-b8-t8192 llama3-8b-4bit-b1-t8192; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r2_run10_bench_$W.json 2> gpurun_out/r2_run10_bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/r2_run10_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3), 'launches', d['gpu_launches']//16, d['clocks']); print({k: v['mean_us'] for k, v in d['kernels'].items()})"
done
