#!/bin/bash
# run 38: cluster norm at every row count (tests) + programmatic dependent launch of the GEMV class (mask 64) and attention (8)
mkdir -p gpurun_out
O=gpurun_out/r2_run38
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layouts.py tests/test_gpu_model.py -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -3
for PM in 0 64 72; do
for W in llama3-8b-4bit-b1-t8192 llama3-8b-4bit-b8-t8192; do
  B200P_PDL_MASK=$PM timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_pm${PM}_$W.json 2> ${O}_bench_pm${PM}_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_pm${PM}_$W.json')); print('PDL_MASK=$PM $W ms/step', round(d['ms_per_step'],3))"
done
done 2>&1 | tee ${O}_pdl_mask_small_batch.txt
