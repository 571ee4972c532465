#!/bin/bash
# run 31: rewritten nucleus sampler (lane-private histograms) + argmax with two loads in flight
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -m gpu -q --timeout 300 -p no:cacheprovider -k "top_p or argmax or sampl or generate_step" 2>&1 | tail -8
timeout 200 python scripts/sampler_bench.py 2>&1 | tee gpurun_out/r31_sampler_bench.txt | tail -14
