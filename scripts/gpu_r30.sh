#!/bin/bash
# run 30: fp32-accumulator limit 64 -> 128 rows, Phi-3 checkpoint test, B=256 point of configs[2] with the context capped at 4096
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -8
run() { timeout 400 python scripts/shard_proxy.py "$@" 2> gpurun_out/r30_err.txt | tee -a gpurun_out/r30_shard_proxy.jsonl | cut -c1-330; tail -3 gpurun_out/r30_err.txt; }
: > gpurun_out/r30_shard_proxy.jsonl
run --config llama-3-70b --world 8 --batch 128 --context 16384
run --config llama-3-70b --world 8 --batch 128 --context 256
for W in llama3-8b-4bit-b256-t4096; do
  timeout 400 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r30_bench_$W.json 2> gpurun_out/r30_bench_$W.err
  tail -3 gpurun_out/r30_bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/r30_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3)); print({k:v for k,v in d['kernels'].items()})"
done
