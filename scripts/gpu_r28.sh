#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -5
python scripts/quantize_bench.py 2>&1 | grep dequantize
for W in mistral-7b-4bit-b8-t256 qwen3-14b-4bit-b64-t256; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3))"
done
