#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -5
for aw in 8 0; do
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b1-t8192 llama3-8b-4bit-b8-t8192 llama3-8b-4bit-b64-t8192; do
  B200P_ATTN_WARPS=$aw timeout 600 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$W.json'))
print('ATTN_WARPS=$aw $W ms/step',round(d['ms_per_step'],3),'attn us',d['kernels']['b200p_attn']['mean_us'])
PY
done
done
