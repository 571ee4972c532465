#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider -x 2>&1 | tail -25
timeout 300 python scripts/qmm_sweep.py 1,4,8,16 2>&1 | tail -24
for W in llama3-8b-4bit-b1-t8192 llama3-8b-4bit-b8-t8192; do
  timeout 600 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$W.json'))
print('$W ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],3),'frac',round(d['roofline_detail']['step_frac_of_hbm_peak'],3))
for k,v in d['kernels'].items(): print('  ',k,v)
PY
done
