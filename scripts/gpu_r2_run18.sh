#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/qmm_layer_bench.py 8,64 llama3-8b-acc 2>&1 | tee gpurun_out/r2_run18_acc_epilogue.txt
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
from mlx_parallm_b200 import ops
for M, I in ((64, 14336), (64, 7168), (64, 1792), (8, 14336)):
    gu = torch.zeros((M, 2 * I), dtype=torch.float32, device='cuda'); out = torch.empty((M, I), dtype=torch.float16, device='cuda')
    ops.silu_mul(gu, out, out_perm=True); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(64): ops.silu_mul(gu, out, out_perm=True)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f'silu_mul M={M} I={I}: {e0.elapsed_time(e1) / 64 * 1e3:.2f} us per launch (graph of 64)')
PY
