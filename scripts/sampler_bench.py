"""Time the sampling head (north star kernel 3): argmax, categorical (top_p = 1) and nucleus sampling on [B, V] float32 logits."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from mlx_parallm_b200 import _lib, ops

_lib.require_device()
g = torch.Generator(device="cuda")
g.manual_seed(3)


def timeit(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3


for V in (32000, 128256, 151936):
    for B in (1, 8, 64, 256):
        # LM-like logits: a few large values over a wide gaussian body
        lg = torch.randn((B, V), generator=g, device="cuda") * 2.5
        lg[:, :50] += 8.0
        out = torch.empty((B,), dtype=torch.int32, device="cuda")
        pr = torch.empty((B,), dtype=torch.float32, device="cuda")
        t_arg = timeit(lambda: ops.argmax(lg, out=out, probs_out=pr))
        t_cat = timeit(lambda: ops.sample_top_p(lg, 0.7, 1.0, 1, 2, out=out, probs_out=pr))
        t_p9 = timeit(lambda: ops.sample_top_p(lg, 0.7, 0.9, 1, 2, out=out, probs_out=pr))
        t_p5 = timeit(lambda: ops.sample_top_p(lg, 1.0, 0.5, 1, 2, out=out, probs_out=pr))
        mb = B * V * 4 / 1e6
        print(f"V={V} B={B} ({mb:.1f} MB fp32): argmax {t_arg:.1f} us | categorical T=0.7 {t_cat:.1f} us | top_p 0.9 T=0.7 {t_p9:.1f} us | "
              f"top_p 0.5 T=1 {t_p5:.1f} us", flush=True)
