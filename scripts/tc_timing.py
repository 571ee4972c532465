"""Where do the roles of qmm_tc_kernel wait?  Needs the timing build (scripts/build_tc_debug.sh t) selected with
B200P_LIB=build/dbgt/libb200parallm.so.  Prints, for CTA 0 of the last launch, the cycles each role spent in each of its
barrier waits next to the role's total.

    B200P_LIB=$PWD/build/dbgt/libb200parallm.so python scripts/tc_timing.py M N K epi [small]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlx_parallm_b200 import _lib, ops

L = _lib.lib()
_lib.require_device()
M, N, K, epi = [int(v) for v in sys.argv[1:5]]
small = len(sys.argv) > 5 and sys.argv[5] == "small"
dt = torch.float16
g = torch.Generator(device="cuda").manual_seed(1)
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
for i in range(8):
    ops.qmm(x, *mats[i % 4], 64, 4, epi, out=out, workspace=ws, x_perm=True, y_perm=epi in (0, 3), small_scale=small)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 32)()
L.b200p_debug_tc_timing.restype = C.c_int
assert L.b200p_debug_tc_timing(buf) == 0
v = list(buf)
names = ["W producer: empty_w", "W producer: sb_empty", "W producer TOTAL", "X producer: empty_x", "X producer TOTAL",
         "MMA: full_x", "MMA: full_a", "MMA: accum_empty", "MMA TOTAL", "dequant w0: full_w", "dequant w0: sb_full",
         "dequant w0: empty_a", "dequant w0 TOTAL", "epilogue: accum_full", "epilogue TOTAL (to end of tile epilogues)",
         "(unused)", "dequant w0 section: LDS of the codes + slot release", "dequant w0 section: dequant math (+ clock reads)",
         "dequant w0 section: tcgen05.st + wait::st", "dequant w0 section: fence + syncwarp + arrive"]
print(f"M={M} N={N} K={K} epi={epi} small={small}: cycles in CTA 0")
for n, c in zip(names, v):
    print(f"  {n:45s} {c:9d}")
