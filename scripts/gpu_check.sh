#!/bin/bash
# One GPU-box pass: tests, smoke, bench.  Outputs land in gpurun_out/ (merged back by gpurun).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python -c "import mlx.core as mx; print('mlx', mx.__version__)" > gpurun_out/mlx_probe.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider 2>&1 | tail -60 > gpurun_out/pytest_gpu.txt
echo "pytest exit: $?" >> gpurun_out/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.txt 2>&1
echo "smoke exit: $?" >> gpurun_out/smoke.txt
timeout 600 python bench.py --workload small --sweep '' --steps 8 > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err
timeout 900 python bench.py ${BENCH_ARGS:---steps 16 --warmup 3} > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit: $?" >> gpurun_out/bench.err
tail -5 gpurun_out/pytest_gpu.txt; cat gpurun_out/smoke.txt | tail -3; cat gpurun_out/bench.json | head -c 3000
