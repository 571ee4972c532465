#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -5
for dbg in 64 0; do
  echo "B200P_TC_DEBUG=$dbg"
  B200P_TC_DEBUG=$dbg timeout 300 python scripts/qmm_sweep.py 16,64 2>&1 | grep -v "^N K" | awk '{print $1,$2,$4,$8,$9}'
  for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b64-t8192; do
  B200P_TC_DEBUG=$dbg timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/bench_dbg.json 2> gpurun_out/bench_dbg.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_dbg.json')); print('TC_DEBUG=$dbg $W ms/step', round(d['ms_per_step'],3))"
  done
done
B200P_TC_DEBUG=64 python scripts/prefill_profile.py 8 256 2>&1 | head -1
python scripts/prefill_profile.py 8 256 2>&1 | head -1
