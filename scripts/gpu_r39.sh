#!/bin/bash
# run 39: split-K merge of the decode attention by the last-arriving CTA (no attn_reduce launch)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -4
for W in llama3-8b-4bit-b64-t8192 llama3-8b-4bit-b64-t256 llama3-8b-4bit-b8-t8192 llama3-8b-4bit-b1-t8192; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r39_bench_$W.json 2> gpurun_out/r39_bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/r39_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3), 'attn us', d['kernels']['b200p_attn'], 'launches', d['gpu_launches']//16, d['clocks'])"
done
