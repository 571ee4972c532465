#!/bin/bash
# the exact launch the driver uses for N > 1, stdout captured separately: it must contain ONE JSON line and nothing else
N=${1:-2}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $RUN bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/driverlike_tp$N.json 2> gpurun_out/driverlike_tp$N.err; echo "ours exit $?"
echo "stdout lines: $(wc -l < gpurun_out/driverlike_tp$N.json)"; python -c "
import json; d=json.loads(open('gpurun_out/driverlike_tp$N.json').read()); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['config'], d['sweep'], d['prefill'], d['gpu_launches'])"
timeout 300 $RUN bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/driverlike_ref_tp$N.json 2> gpurun_out/driverlike_ref_tp$N.err; echo "reference exit $?"
echo "stdout lines: $(wc -l < gpurun_out/driverlike_ref_tp$N.json)"; head -c 300 gpurun_out/driverlike_ref_tp$N.json; echo
