"""Launch the tensor-path GEMM a few times eagerly (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlx_parallm_b200 import ops, _lib
L = _lib.lib(); _lib.require_device()
M, N, K, epi = [int(v) for v in sys.argv[1:5]]
small = len(sys.argv) > 5 and sys.argv[5] == "small"
dt = torch.float16
g = torch.Generator(device="cuda"); g.manual_seed(1)
mats = []
for i in range(4):
    w = torch.randint(-2**31, 2**31, (N, K // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    s = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
    b = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
    mats.append((w, s, b))
x = torch.randn((M, K), device="cuda").to(dt)
ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
out = torch.zeros((M, N if epi != 3 else N // 2), device="cuda", dtype=torch.float32 if epi in (1, 2) else dt)
_lib.check(L.b200p_qmm_set_path(2))
evs = []
for i in range(8):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.qmm(x, *mats[i % 4], 64, 4, epi, out=out, workspace=ws, x_perm=True, y_perm=epi in (0, 3), small_scale=small)
    e1.record()
    evs.append((e0, e1))
torch.cuda.synchronize()
print("event us per launch:", [round(a.elapsed_time(b) * 1e3, 1) for a, b in evs])
print("done")
