#!/bin/bash
# exchange kernels on one device (simulated ranks) + small-batch kernel breakdowns + PDL modes
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_exchange.py -m gpu -q --timeout 120 -p no:cacheprovider 2>&1 | tail -15
for W in llama3-8b-4bit-b1-t8192 llama3-8b-4bit-b8-t8192; do
  timeout 600 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$W.json'))
print('$W ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],3),'frac',round(d['roofline_detail']['step_frac_of_hbm_peak'],3))
for k,v in d['kernels'].items(): print('  ',k,v)
PY
done
for m in 0 2 1; do
  for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b1-t8192; do
  B200P_PDL=$m timeout 600 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pdl${m}_$W.json 2> gpurun_out/bench_pdl${m}_$W.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_pdl${m}_$W.json')); print('PDL=$m $W ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"
  done
done
