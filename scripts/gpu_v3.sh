#!/bin/bash
# tensor-path GEMM: correctness (tc_check + pytest), timing, bisect, short bench
mkdir -p gpurun_out
timeout 300 python scripts/tc_check.py > gpurun_out/tc_check.txt 2>&1; echo "tc_check exit $?" >> gpurun_out/tc_check.txt
tail -12 gpurun_out/tc_check.txt
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -25 > gpurun_out/pytest_gpu.txt
cat gpurun_out/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 600 python bench.py --workload llama3-8b-4bit-b64-t256 --sweep= --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_t256.json 2> gpurun_out/bench_t256.err; tail -3 gpurun_out/bench_t256.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_t256.json'))
print('ms/step',d['ms_per_step'],'frac',d['roofline_detail']['step_frac_of_hbm_peak'])
for k,v in d['kernels'].items(): print(' ',k,v)
PY
