#!/bin/bash
# run 36: auto rule with the mma.sync small-matrix GEMM at 5..16 rows: GEMM tests + the sweep points it touches
mkdir -p gpurun_out
O=gpurun_out/r2_run36
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_layouts.py -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -3
for P in 4 16; do
for W in llama3-8b-4bit-b8-t8192 mistral-7b-4bit-b8-t256; do
  B200P_MMA_MAX_M=$P timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_m${P}_$W.json 2> ${O}_bench_m${P}_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_m${P}_$W.json')); print('MMA_MAX_M=$P $W ms/step', round(d['ms_per_step'],3))"
done
done 2>&1 | tee ${O}_mma_small_gemm_steps.txt
