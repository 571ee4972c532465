#!/bin/bash
# run 29: per-GPU share of the TP configs on one GPU (no exchange): configs[4] (70B-class 8-bit) and configs[3] (Qwen3-14B)
mkdir -p gpurun_out
run() { timeout 400 python scripts/shard_proxy.py "$@" 2> gpurun_out/r29_err.txt | tee -a gpurun_out/r29_shard_proxy.jsonl | cut -c1-400; tail -3 gpurun_out/r29_err.txt; }
: > gpurun_out/r29_shard_proxy.jsonl
run --config llama-3-70b --world 8 --batch 128 --context 16384
run --config llama-3-70b --world 8 --batch 64 --context 32768
run --config llama-3-70b --world 8 --batch 128 --context 256
run --config qwen3-14b --world 8 --batch 64 --context 256
run --config qwen3-14b --world 4 --batch 64 --context 256
run --config qwen3-14b --world 2 --batch 64 --context 256
