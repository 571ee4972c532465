#!/bin/bash
# round-end pass 2: smoke, the GPU suite, the default bench line, the reference arm, the ncu launch list of the bench command
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/final2_pytest_gpu.txt 2>&1; tail -2 gpurun_out/final2_pytest_gpu.txt
timeout 900 python bench.py > gpurun_out/final2_bench.json 2> gpurun_out/final2_bench.err; echo "bench exit $? lines $(wc -l < gpurun_out/final2_bench.json)"
python -c "
import json; d=json.load(open('gpurun_out/final2_bench.json')); print(d['ms_per_step'], d['value'], d['e2e'], d['roofline'], d['cpu_baseline'], d['sweep'], d['prefill'], d['clocks'], d['gpu_launches'])"
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/final2_bench_reference.json 2> gpurun_out/final2_bench_reference.err; echo "reference exit $?"; head -c 400 gpurun_out/final2_bench_reference.json; echo
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/final2_ncu_launches_headline.csv python bench.py --steps 2 --warmup 3 --sweep= --no-cpu-baseline --no-prefill > gpurun_out/final2_ncu.log 2>&1; echo "ncu exit $?"; wc -l gpurun_out/final2_ncu_launches_headline.csv
