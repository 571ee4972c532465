#!/bin/bash
# run 37: attn_reduce with parallel loads; the early (cluster) RMSNorm below 16 rows
mkdir -p gpurun_out
O=gpurun_out/r2_run37
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py tests/test_gpu_model.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "attn or attention or static or greedy or logits" 2>&1 | tail -3
for E in auto 1; do
for W in llama3-8b-4bit-b1-t8192 llama3-8b-4bit-b8-t8192; do
  if [ "$E" = "auto" ]; then unset B200P_RMS_EARLY; else export B200P_RMS_EARLY=$E; fi
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_e${E}_$W.json 2> ${O}_bench_e${E}_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_e${E}_$W.json')); print('RMS_EARLY=$E $W ms/step', round(d['ms_per_step'],3), {k: round(v['mean_us'],1) for k, v in d['kernels'].items() if 'attn' in k or 'rms' in k})"
done
done 2>&1 | tee ${O}_small_batch_steps.txt
unset B200P_RMS_EARLY
timeout 300 python scripts/step_timeline.py llama3-8b-4bit-b1-t8192 ${O}_timeline_b1_t8192 > /dev/null 2> ${O}_timeline.err
sed -n 1,14p ${O}_timeline_b1_t8192.txt
