"""Where does the tensor-path GEMM spend its time?  Times a few decode shapes with parts of the kernel
switched off (B200P_TC_DEBUG bits) and with L2-resident vs HBM-streamed weights."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlx_parallm_b200 import ops, _lib
L = _lib.lib(); _lib.require_device()

def make(N, K, dt, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    w = torch.randint(-2**31, 2**31, (N, K // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    s = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
    b = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
    return w, s, b

def bench(M, N, K, epi, rotate, reps=20):
    dt = torch.float16
    nm = max(2, int(300e6 // (N * K / 2)) + 1) if rotate else 1
    mats = [make(N, K, dt, i) for i in range(nm)]
    x = torch.randn((M, K), device="cuda").to(dt)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    out = torch.zeros((M, N if epi != 3 else N // 2), device="cuda", dtype=torch.float32 if epi in (1, 2) else dt)
    _lib.check(L.b200p_qmm_set_path(2))
    for i in range(3): ops.qmm(x, *mats[i % nm], 64, 4, epi, out=out, workspace=ws)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps): ops.qmm(x, *mats[i % nm], 64, 4, epi, out=out, workspace=ws)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    L.b200p_qmm_set_path(0)
    ws.zero_()
    return e0.elapsed_time(e1) / reps * 1e3

shapes = [(64, 28672, 4096, 2), (64, 4096, 14336, 2), (64, 4096, 4096, 2), (16, 28672, 4096, 2)]
if len(sys.argv) > 1 and sys.argv[1] == "all":
    shapes += [(64, 4096, 4096, 2), (64, 4096, 14336, 2), (64, 6144, 4096, 0)]
# bits: 2 no MMA, 4 no epilogue, 8 no TMEM store of A, 16 no LDS + dequant math
# bits: 1 no X loads, 2 no MMA issue, 4 no epilogue stores, 8 no A stores
modes = [("full", 0), ("-sb", 32), ("-x-mma-epi", 7), ("-x-mma-epi-sb", 39), ("-all", 47)]
for (M, N, K, epi) in shapes:
    byts = N * K / 2 + 2 * N * (K // 64) * 2
    row = []
    for name, dbg in modes:
        os.environ["B200P_TC_DEBUG"] = str(dbg)
        row.append(f"{name}={bench(M, N, K, epi, True):.1f}")
    os.environ["B200P_TC_DEBUG"] = "0"
    print(f"M={M} N={N} K={K} epi{epi} ideal={byts/6.55e6:.1f}us : " + "  ".join(row), flush=True)
