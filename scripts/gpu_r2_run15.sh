#!/bin/bash
# round 2, run 15: dequant warps help drain the last whole tile
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_fullsize.py -x -q --timeout 300 -p no:cacheprovider 2>&1 | tail -5
timeout 300 python scripts/qmm_layer_bench.py 1,8,64 llama3-8b gateup,head,qkv 2>&1 | tee gpurun_out/r2_run15_qmm_layer_bench.txt
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b64-t8192; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r2_run15_bench_$W.json 2> gpurun_out/r2_run15_bench_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/r2_run15_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3))"
done
