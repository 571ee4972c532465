#!/bin/bash
# run 35: mma.sync small-matrix GEMM (qmm_mma.cu): tests, then per-matrix times against the tile kernels (path 2)
mkdir -p gpurun_out
O=gpurun_out/r2_run35
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "qmm" 2>&1 | tail -5
for P in 2 4; do
  QMM_BENCH_PATH=$P timeout 300 python scripts/qmm_layer_bench.py 8,16,32,64 small-acc 2>&1
done | tee ${O}_small_gemm_paths.txt
