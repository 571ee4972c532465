#!/bin/bash
# run 41: the early norms and the small-matrix GEMM follow the global PDL switch; tests with PDL off, steps with PDL off / on
mkdir -p gpurun_out
O=gpurun_out/r2_run41
B200P_PDL=0 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -2
for P in 0 2; do
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b8-t8192; do
  B200P_PDL=$P timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_pdl${P}_$W.json 2> ${O}_bench_pdl${P}_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_pdl${P}_$W.json')); print('B200P_PDL=$P $W ms/step', round(d['ms_per_step'],3))"
done
done 2>&1 | tee ${O}_pdl_switch.txt
