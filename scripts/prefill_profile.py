"""Per-kernel CUDA-event breakdown of one eager prefill pass (Llama-3-8B 4-bit, [B, Lp] prompt batch)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mlx_parallm_b200 import ops, _lib
_lib.require_device()
B, Lp = int(sys.argv[1]), int(sys.argv[2])
cfg, model = bench.build_model("llama-3-8b", "cuda:0")
ids = torch.randint(0, cfg["vocab_size"], (B, Lp), device="cuda")
cache = model.make_cache(B, step=Lp + 8)
for rep in range(2):
    for c in cache:
        c.reset(B)
    torch.cuda.synchronize()
    ops.PROFILER = {} if rep == 1 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    model(ids, cache=cache, last_only=True)
    e1.record()
    torch.cuda.synchronize()
prof, ops.PROFILER = ops.PROFILER, None
tot = 0.0
rows = []
for k, evs in prof.items():
    t = sum(s.elapsed_time(e) for s, e in evs)
    rows.append((t, k, len(evs)))
    tot += t
print(f"prefill B={B} Lp={Lp}: {e0.elapsed_time(e1):.2f} ms wall (instrumented), {B * Lp / e0.elapsed_time(e1) * 1e3:.0f} tok/s")
for t, k, n in sorted(rows, reverse=True):
    print(f"  {k:34s} calls {n:4d}  total {t:8.2f} ms  mean {t / n * 1e3:9.1f} us  share {100 * t / tot:5.1f} %")
