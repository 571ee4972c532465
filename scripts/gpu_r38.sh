#!/bin/bash
# run 38: rope + KV append, one warp per (token, head), no loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -4
for W in llama3-8b-4bit-b64-t256 qwen3-14b-4bit-b64-t256 llama3-8b-4bit-b64-t8192; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r38_bench_$W.json 2> gpurun_out/r38_bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/r38_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3), 'rope us', d['kernels']['b200p_rope_kv_append'], d['clocks'])"
done
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__cycles_active.avg --clock-control none -k regex:"rope" --launch-skip 100 --launch-count 4 --csv --log-file gpurun_out/r38_rope.csv python bench.py --workload llama3-8b-4bit-b64-t256 --sweep= --steps 4 --warmup 3 --no-cpu-baseline --no-prefill > /dev/null 2>&1
grep -E "gpu__time_duration|cycles_active" gpurun_out/r38_rope.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | head -8
