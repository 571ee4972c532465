#!/bin/bash
# run 39: gate|up with the fused silu*mul epilogue through the mma.sync GEMV at M <= 4 (B200P_GEMV_SILU_MAX_M) against the tile kernel
mkdir -p gpurun_out
O=gpurun_out/r2_run39
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layouts.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "qmv or qmm" 2>&1 | tail -3
for S in 0 4; do
  B200P_GEMV_SILU_MAX_M=$S timeout 300 python bench.py --workload llama3-8b-4bit-b1-t8192 --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_s${S}.json 2> ${O}_bench_s${S}.err
  python -c "
import json; d=json.load(open('${O}_bench_s${S}.json')); print('GEMV_SILU_MAX_M=$S b1-t8192 ms/step', round(d['ms_per_step'],3), {k: round(v['mean_us'],1) for k, v in d['kernels'].items() if '28672' in k})"
done 2>&1 | tee ${O}_gemv_silu.txt
