#!/bin/bash
# Tensor-parallel check on N GPUs of one box: NCCL bring-up, TP parity test, bench at TP=N.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ -z "$SKIP_CHECK" ]; then
  timeout 300 $RUN scripts/tp_check.py > gpurun_out/tp_check_$N.txt 2>&1; echo "tp_check exit $?" >> gpurun_out/tp_check_$N.txt
  tail -15 gpurun_out/tp_check_$N.txt
fi
if [ -n "$XBENCH" ]; then
  timeout 120 $RUN scripts/tp_exchange_bench.py > gpurun_out/tp_exchange_$N.txt 2>&1; echo "exchange bench exit $?" >> gpurun_out/tp_exchange_$N.txt
  grep -E "^tp|exit|Error|error" gpurun_out/tp_exchange_$N.txt | tail -12
fi
for W in ${WORKLOADS-llama3-8b-4bit-b64-t256} ${EXTRA_WORKLOADS}; do
  timeout 300 $RUN bench.py --gpus $N --steps 16 --warmup 3 --workload $W --sweep= --no-cpu-baseline \
    > gpurun_out/bench_tp${N}_$W.json 2> gpurun_out/bench_tp${N}_$W.err
  echo "bench $W exit $?"; tail -3 gpurun_out/bench_tp${N}_$W.err; head -c 1500 gpurun_out/bench_tp${N}_$W.json; echo
done
