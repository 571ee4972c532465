"""Tensor-path (tcgen05) GEMM vs the CUDA-core GEMV path and a float64 reference; prints errors and GB/s."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mlx_parallm_b200 import ops, _lib

L = _lib.lib(); _lib.require_device()
torch.manual_seed(0)

def make(N, K, bits, dt):
    g = torch.Generator(device="cuda"); g.manual_seed(N * 31 + K)
    w = torch.randint(-2**31, 2**31, (N, K * bits // 32), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    nb = (1 << bits) - 1
    s = ((torch.rand((N, K // 64), generator=g, device="cuda") + 0.5) * (2 / nb) / K ** 0.5)
    s = (s * torch.where(torch.rand((N, K // 64), generator=g, device="cuda") < 0.5, -1.0, 1.0)).to(dt)
    b = (-s.float() * nb / 2 + torch.randn((N, K // 64), generator=g, device="cuda") * 0.01 / K ** 0.5).to(dt)
    return w, s, b

def run(M, N, K, bits, dt, epi, path):
    w, s, b = make(N, K, bits, dt)
    g = torch.Generator(device="cuda"); g.manual_seed(M)
    x = torch.randn((M, K), generator=g, device="cuda").to(dt)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    res = torch.randn((M, N), generator=g, device="cuda") if epi == 2 else None
    _lib.check(L.b200p_qmm_set_path(path))
    out = res.clone() if epi == 2 else None
    y = ops.qmm(x, w, s, b, 64, bits, epi, out=out, workspace=ws)
    torch.cuda.synchronize()
    L.b200p_qmm_set_path(0)
    return y.float(), (x, w, s, b, res)

def ref64(x, w, s, b, bits, epi, res):
    wd = ops.dequantize(w, s.float(), b.float(), 64, bits, mode=1).double()     # exact q*s+b in f32 -> f64
    acc = x.double() @ wd.t()
    if epi == 2: acc = acc + res.double()
    if epi == 3:
        n2 = acc.shape[1] // 2
        acc = torch.nn.functional.silu(acc[:, :n2]) * acc[:, n2:]
    return acc

cases = [(8, 256, 256, 4, torch.float16, 1), (16, 512, 1024, 4, torch.float16, 1), (64, 4096, 4096, 4, torch.float16, 1),
         (64, 6144, 4096, 4, torch.float16, 0), (64, 4096, 4096, 4, torch.float16, 2), (64, 28672, 4096, 4, torch.float16, 3),
         (64, 4096, 14336, 4, torch.float16, 2), (5, 1000, 192, 4, torch.float16, 1), (100, 16032, 4096, 4, torch.bfloat16, 1),
         (300, 1024, 512, 4, torch.float16, 0), (64, 2048, 2048, 8, torch.float16, 1), (33, 7168, 5120, 4, torch.bfloat16, 0),
         (256, 128256, 4096, 4, torch.float16, 1)]
only = sys.argv[1:] 
bad = 0
for (M, N, K, bits, dt, epi) in ([] if '--nocheck' in only else cases):
    try:
        y2, (x, w, s, b, res) = run(M, N, K, bits, dt, epi, 2)
        r = ref64(x, w, s, b, bits, epi, res)
        e2 = ((y2.double() - r).abs().max() / r.abs().max()).item()
        y1, _ = run(M, N, K, bits, dt, epi, 1) if M <= 64 else (None, None)
        e1 = ((y1.double() - r).abs().max() / r.abs().max()).item() if y1 is not None else float("nan")
        ok = e2 < (4e-3 if dt == torch.float16 else 2e-2)
        bad += (not ok)
        print(f"M={M:4d} N={N:6d} K={K:5d} b{bits} {str(dt)[6:]:9s} epi{epi}: tensor err {e2:.2e}  gemv err {e1:.2e}  {'OK' if ok else 'FAIL'}", flush=True)
    except Exception as ex:
        bad += 1
        print(f"M={M} N={N} K={K} b{bits} epi{epi}: EXC {ex}", flush=True)

# timing at the decode shapes of Llama-3-8B, B=64 (weights do not fit L2 across the 4 different matrices + rotation)
def bench(M, N, K, bits, dt, epi, path, reps=20):
    mats = [make(N + 0, K, bits, dt) for _ in range(max(2, int(300e6 // (N * K * bits / 8)) + 1))]   # > 126 MB L2 rotation
    x = torch.randn((M, K), device="cuda").to(dt)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    out = torch.zeros((M, N if epi != 3 else N // 2), device="cuda", dtype=torch.float32 if epi in (1, 2) else dt)
    _lib.check(L.b200p_qmm_set_path(path))
    for i in range(3): ops.qmm(x, *mats[i % len(mats)], 64, bits, epi, out=out, workspace=ws)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps): ops.qmm(x, *mats[i % len(mats)], 64, bits, epi, out=out, workspace=ws)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record(); torch.cuda.synchronize()
    L.b200p_qmm_set_path(0)
    ms = e0.elapsed_time(e1) / reps
    byts = N * K * bits / 8 + 2 * N * (K // 64) * 2
    return ms, byts / ms / 1e6

if "--notime" not in only:
    for (M, N, K, epi) in [(64, 6144, 4096, 0), (64, 4096, 4096, 2), (64, 28672, 4096, 3), (64, 4096, 14336, 2), (64, 128256, 4096, 1),
                           (8, 28672, 4096, 3), (16, 28672, 4096, 3), (128, 28672, 4096, 3), (256, 28672, 4096, 3)]:
        ms, gbs = bench(M, N, K, 4, torch.float16, epi, 2)
        print(f"time M={M:3d} N={N:6d} K={K:5d} epi{epi}: {ms*1e3:8.1f} us  {gbs:7.0f} GB/s ({gbs/6553*100:.0f}% of 6553)", flush=True)
print("FAILURES:", bad)
