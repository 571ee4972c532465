#!/bin/bash
# round 2: 8-GPU evidence -- exchange soak, headline + weights-bound point at TP=8, configs[3] (Qwen3-14B TP=4/8), configs[4] (70B-class 8-bit TP=8)
mkdir -p gpurun_out
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1].split('/')[-1], '| ms/step', round(d['ms_per_step'], 3), 'tok/s', round(d['value']), 'e2e', round(d['e2e']['value']),
          '| frac of HBM peak', round(d['roofline_detail']['step_frac_of_hbm_peak'], 3), '| exchange', d['config']['tp_exchange'],
          'share', round(d['roofline_detail']['exchange_share_of_step_time'], 3), '| selfcheck', d.get('tp_selfcheck'), '| sweep', d['sweep'], '| clocks', d['clocks'])
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
}
timeout 400 $TR8 scripts/tp_soak.py 102400 64 4096 2>&1 | grep "^tp" | tee gpurun_out/r2_tp8_soak.txt
timeout 500 $TR8 bench.py --gpus 8 --steps 20 --warmup 3 --sweep=llama3-8b-4bit-b64-t256 --no-prefill > gpurun_out/r2_tp8_bench_headline.json 2> gpurun_out/r2_tp8_bench_headline.err
show gpurun_out/r2_tp8_bench_headline.json
timeout 500 $TR8 bench.py --gpus 8 --workload qwen3-14b-4bit-b64-t256 --steps 20 --warmup 3 --sweep= --no-prefill > gpurun_out/r2_tp8_bench_qwen3.json 2> gpurun_out/r2_tp8_bench_qwen3.err
show gpurun_out/r2_tp8_bench_qwen3.json
timeout 500 $TR4 bench.py --gpus 4 --workload qwen3-14b-4bit-b64-t256 --steps 20 --warmup 3 --sweep= --no-prefill > gpurun_out/r2_tp4_bench_qwen3.json 2> gpurun_out/r2_tp4_bench_qwen3.err
show gpurun_out/r2_tp4_bench_qwen3.json
timeout 900 $TR8 bench.py --gpus 8 --workload llama3-70b-8bit-b128-t24576 --steps 8 --warmup 3 --sweep=llama3-70b-8bit-b128-t256 --no-prefill > gpurun_out/r2_tp8_bench_70b.json 2> gpurun_out/r2_tp8_bench_70b.err
show gpurun_out/r2_tp8_bench_70b.json
tail -3 gpurun_out/r2_tp8_bench_70b.err
