#!/bin/bash
# round 2, run 3/4: rotated stage order; plan variants
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -k "qmm" -x -q --timeout 120 -p no:cacheprovider 2>&1 | tail -5
echo "== layer bench (default plan)"
timeout 300 python scripts/qmm_layer_bench.py 1,8,64 llama3-8b 2>&1 | tee gpurun_out/r2_run4_qmm_layer_bench.txt
for V in "B200P_TC_ROT=0" "B200P_TC_R=128" "B200P_TC_R=128 B200P_TC_S=4"; do
  echo "== $V"
  env $V timeout 300 python scripts/qmm_layer_bench.py 64 llama3-8b 2>&1 | tail -5
done
