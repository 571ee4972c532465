#!/bin/bash
# round 2, run 11: A/B of the round-1 GEMM library against today's, same scripts, same box
mkdir -p gpurun_out
for V in r1 cur; do
  if [ $V = r1 ]; then export B200P_LIB=$PWD/build/r1/libb200parallm.so B200P_NO_SMALLSCALE=1; else unset B200P_LIB B200P_NO_SMALLSCALE; fi
  echo "=== $V"
  timeout 300 python scripts/qmm_layer_bench.py 8,64 llama3-8b 2>&1 | grep -v "^model"
  for W in llama3-8b-4bit-b64-t256; do
    timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > gpurun_out/r2_run11_${V}_$W.json 2> gpurun_out/r2_run11_${V}_$W.err
    python -c "
import json; d=json.load(open('gpurun_out/r2_run11_${V}_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'launches', d['gpu_launches']//16); print({k: v['mean_us'] for k, v in d['kernels'].items()})"
  done
done
