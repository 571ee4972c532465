#!/bin/bash
# in-graph timeline of the weights-bound step (CUPTI via torch.profiler) + the bench number of the same step without a profiler
mkdir -p gpurun_out
timeout 300 python bench.py --workload llama3-8b-4bit-b64-t256 --sweep= --steps 32 --warmup 4 --no-cpu-baseline --no-prefill > gpurun_out/r2_run20_bench_b64_t256.json 2> gpurun_out/r2_run20_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r2_run20_bench_b64_t256.json')); print('b64-t256 ms/step', round(d['ms_per_step'],3))"
timeout 300 python scripts/step_timeline.py llama3-8b-4bit-b64-t256 gpurun_out/r2_run20_timeline_b64_t256 > /dev/null 2> gpurun_out/r2_run20_timeline.err
tail -3 gpurun_out/r2_run20_timeline.err
head -30 gpurun_out/r2_run20_timeline_b64_t256.txt
timeout 300 python scripts/step_timeline.py llama3-8b-4bit-b1-t8192 gpurun_out/r2_run20_timeline_b1_t8192 > /dev/null 2>> gpurun_out/r2_run20_timeline.err
head -20 gpurun_out/r2_run20_timeline_b1_t8192.txt
