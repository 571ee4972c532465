"""b200p_quantize / b200p_dequantize throughput at Llama-3-8B matrix shapes (HBM-bound byte work)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlx_parallm_b200 import ops, _lib
_lib.require_device()
for N, K in [(4096, 14336), (28672, 4096), (128256, 4096)]:
    w = (torch.randn((N, K), device="cuda") * 0.02).to(torch.float16)
    for name, fn, byts in (("quantize", lambda: ops.quantize(w, 64, 4), N * K * 2 + N * K // 2 + 4 * N * K // 64),):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"{name} {N}x{K} f16 -> 4-bit g64: {ms * 1e3:.1f} us, {byts / ms / 1e6:.0f} GB/s of algorithmic bytes")
    wq, s, b = ops.quantize(w, 64, 4)
    out = torch.empty_like(w)
    for _ in range(3):
        ops.dequantize(wq, s, b, 64, 4, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ops.dequantize(wq, s, b, 64, 4, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    byts = N * K * 2 + N * K // 2 + 4 * N * K // 64
    print(f"dequantize {N}x{K} 4-bit g64 -> f16: {ms * 1e3:.1f} us, {byts / ms / 1e6:.0f} GB/s of algorithmic bytes")
