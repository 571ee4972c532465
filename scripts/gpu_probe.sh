#!/bin/bash
# HBM / L2 streaming-pattern probe + the regular test / smoke / bench pass
mkdir -p gpurun_out
timeout 120 ./scripts/probe_stream > gpurun_out/probe_stream.txt 2>&1
echo "probe exit $?" >> gpurun_out/probe_stream.txt
cat gpurun_out/probe_stream.txt
bash scripts/gpu_round.sh
