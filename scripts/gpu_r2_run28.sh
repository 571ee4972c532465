#!/bin/bash
# run 28: two MMA-issuing warps + dequant warps publish a k-block only after computing the next (TMEM store latency hidden)
mkdir -p gpurun_out
O=gpurun_out/r2_run28
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "qmm or tc or gemm or epi" 2>&1 | tail -4
echo "== layer bench"
timeout 300 python scripts/qmm_layer_bench.py 1,8,64 llama3-8b 2>&1 | tee ${O}_qmm_layer_bench.txt
for W in llama3-8b-4bit-b64-t256 llama3-8b-4bit-b8-t8192 llama3-8b-4bit-b1-t8192; do
  timeout 300 python bench.py --workload $W --sweep= --steps 16 --warmup 3 --no-cpu-baseline --no-prefill > ${O}_bench_$W.json 2> ${O}_bench_$W.err
  python -c "
import json; d=json.load(open('${O}_bench_$W.json')); print('$W ms/step', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'frac', round(d['roofline_detail']['step_frac_of_hbm_peak'],3))"
done
export B200P_LIB=$PWD/build/dbgt/libb200parallm.so
export B200P_PDL=0
for a in "8 28672 4096 3" "64 28672 4096 3"; do
  python scripts/tc_timing.py $a small
done 2>&1 | tee ${O}_tc_timing_nopdl.txt
