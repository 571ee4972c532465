#!/bin/bash
mkdir -p gpurun_out
for m in 0 2 1; do
  B200P_PDL=$m timeout 600 python bench.py --workload llama3-8b-4bit-b64-t256 --sweep= --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pdl$m.json 2> gpurun_out/bench_pdl$m.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_pdl$m.json')); print('PDL=$m ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"
done
