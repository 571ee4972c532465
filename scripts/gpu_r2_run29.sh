#!/bin/bash
# run 29: "early" RMSNorm (128 threads x 32 registers, PDL attribute, dependents released before its own wait) so that the GEMM
# after a norm is scheduled while the GEMM before it drains.  B200P_RMS_EARLY = 0 / 1 forces it off / on; unset = the auto rule.
mkdir -p gpurun_out
O=gpurun_out/r2_run29
for E in 0 1 auto; do
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b1-t8192; do
  if [ "$E" = "auto" ]; then unset B200P_RMS_EARLY; else export B200P_RMS_EARLY=$E; fi
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_e${E}_$W.json 2> ${O}_bench_e${E}_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_e${E}_$W.json')); print('RMS_EARLY=$E $W ms/step', round(d['ms_per_step'],3))"
done
done 2>&1 | tee ${O}_rms_early.txt
unset B200P_RMS_EARLY
timeout 300 python scripts/step_timeline.py llama3-8b-4bit-b64-t256 ${O}_timeline_early_b64_t256 > /dev/null 2> ${O}_timeline.err
head -14 ${O}_timeline_early_b64_t256.txt
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_kernels.py -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -3
