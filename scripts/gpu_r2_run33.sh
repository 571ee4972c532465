#!/bin/bash
# run 33: one-pass cluster RMSNorm (B200P_RMS_CLUSTER=0 disables it: the two-pass early kernel of run 29)
mkdir -p gpurun_out
O=gpurun_out/r2_run33
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "rmsnorm" 2>&1 | tail -4
for C in 0 1; do
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b64-t8192; do
  B200P_RMS_CLUSTER=$C timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_c${C}_$W.json 2> ${O}_bench_c${C}_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_c${C}_$W.json')); print('RMS_CLUSTER=$C $W ms/step', round(d['ms_per_step'],3))"
done
done 2>&1 | tee ${O}_rms_cluster.txt
timeout 300 python scripts/step_timeline.py llama3-8b-4bit-b64-t256 ${O}_timeline_b64_t256 > /dev/null 2> ${O}_timeline.err
sed -n 1,12p ${O}_timeline_b64_t256.txt
