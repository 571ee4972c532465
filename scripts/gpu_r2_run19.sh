#!/bin/bash
for i in 1 2 3 4 5; do
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_frontends.py tests/test_gpu_lora.py -x -q --timeout 300 -p no:cacheprovider 2>&1 | grep -E "^E  |^tests/|passed|failed|Error" | head -30
done
