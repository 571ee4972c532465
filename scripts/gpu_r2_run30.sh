#!/bin/bash
# run 30: b200p_attn_decode_static (no-allocation decode attention): its test alone in a fresh process, then the whole GPU
# suite with it collected last
mkdir -p gpurun_out
O=gpurun_out/r2_run30
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 120 -p no:cacheprovider 2>&1 | tail -6 | tee ${O}_no_alloc_alone.txt
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider 2>&1 | tail -8 | tee ${O}_pytest_gpu.txt
