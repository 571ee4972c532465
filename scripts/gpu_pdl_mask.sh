#!/bin/bash
# which non-GEMM kernel classes profit from programmatic dependent launch (B200P_PDL_MASK bits: 2 rmsnorm, 4 rope, 8 attn, 16 silu, 32 misc)
mkdir -p gpurun_out
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b1-t8192; do
for m in 0 2 4 8 16 32 18 22 30 62; do
  B200P_PDL_MASK=$m timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/bench_mask.json 2> gpurun_out/bench_mask.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_mask.json')); print('PDL_MASK=$m $W ms/step', round(d['ms_per_step'],3))"
done
done
