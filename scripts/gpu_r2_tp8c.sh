#!/bin/bash
# round 2, final build: TP=8 bench lines (headline + the weights-bound point, Qwen3-14B) with the self-check against the unsharded model
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --steps 16 --warmup 3 --sweep=llama3-8b-4bit-b64-t256 --no-prefill > gpurun_out/r2_tp8_final_bench_headline.json 2> gpurun_out/r2_tp8_final_bench_headline.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_tp8_final_bench_headline.json'))
print('tp8 headline ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'selfcheck', d.get('tp_selfcheck'), 'exchange share', round(d['roofline_detail']['exchange_share_of_step_time'],3), 'sweep', d['sweep'])
PY
tail -3 gpurun_out/r2_tp8_final_bench_headline.err
timeout 600 $TR bench.py --gpus 8 --workload qwen3-14b-4bit-b64-t256 --steps 16 --warmup 3 --sweep= --no-prefill > gpurun_out/r2_tp8_final_bench_qwen3.json 2> gpurun_out/r2_tp8_final_bench_qwen3.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_tp8_final_bench_qwen3.json'))
print('tp8 qwen3-14b b64 t256 ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'selfcheck', d.get('tp_selfcheck'), 'exchange share', round(d['roofline_detail']['exchange_share_of_step_time'],3))
PY
