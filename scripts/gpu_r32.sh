#!/bin/bash
# run 32: argmax / nucleus sampler split over thread-block clusters (DSMEM all-to-all)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -8
for S in 1 2 4 8; do echo "--- forced cluster $S"; B200P_SAMPLER_CLUSTER=$S timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q --timeout 300 -p no:cacheprovider -k "top_p or argmax" 2>&1 | tail -3; done
timeout 200 python scripts/sampler_bench.py 2>&1 | tee gpurun_out/r32_sampler_bench.txt | tail -14
