"""GEMV (CUDA cores) vs tensor path (tcgen05) of b200p_qmm over M: the evidence behind the crossover constant MIN_TC_M.

For each layer shape of Llama-3-8B and each M, both paths are timed in a CUDA graph of 4 x 8 launches rotating over 4
weight copies (> L2), device events, and reported as us per launch and GB/s of algorithmic weight bytes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlx_parallm_b200 import _lib, ops

L = _lib.lib()
_lib.require_device()
_lib.check(L.b200p_qmm_set_min_tensor_m(1))
dt = torch.float16
g = torch.Generator(device="cuda").manual_seed(1)
shapes = [(6144, 4096, 2), (4096, 4096, 2), (28672, 4096, 2), (4096, 14336, 2), (128256, 4096, 1)]
Ms = [1, 2, 4, 8, 16, 64]
if len(sys.argv) > 1:
    Ms = [int(v) for v in sys.argv[1].split(",")]
print("N K epi M : cuda-core gemv us (GB/s) | tcgen05 tiles us (GB/s) | mma.sync gemv us (GB/s)")
for N, K, epi in shapes:
    ncopy = 4 if N * K < 2e8 else 2
    mats = []
    for i in range(ncopy):
        w = torch.randint(-2**31, 2**31, (N, K // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        s = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
        b = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
        mats.append((w, s, b))
    wbytes = N * K // 2 + 2 * N * (K // 64) * 2
    for M in Ms:
        x = torch.randn((M, K), device="cuda").to(dt)
        ws = torch.zeros((max(ops.qmm_workspace_bytes(max(M, 8), N, K), 16),), dtype=torch.uint8, device="cuda")
        out = torch.zeros((M, N), device="cuda", dtype=torch.float32)
        res = []
        for path in (1, 2, 3):
            _lib.check(L.b200p_qmm_set_path(path))
            if (path == 2 and not L.b200p_qmm_tensor_supported(M, N, K, 64, 4, 1, epi)) or (path == 3 and M > 16) or (path == 1 and M > 16):
                res.append(float("nan"))
                continue
            reps = 8

            def body():
                for r in range(reps):
                    for m in mats:
                        ops.qmm(x, *m, 64, 4, epi, out=out, workspace=ws)
            body()
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                body()
            gr.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                gr.replay()
            e1.record()
            torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1) * 1e3 / (3 * reps * len(mats)))
            del gr
        _lib.check(L.b200p_qmm_set_path(0))
        print(f"{N} {K} {epi} {M} : " + " | ".join(f"{r:.1f} ({wbytes / r / 1e3:.0f})" for r in res), flush=True)
    del mats
    torch.cuda.empty_cache()
