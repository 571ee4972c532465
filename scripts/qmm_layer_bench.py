"""Per-launch time of the tensor-path GEMM on the layer shapes of a model (default Llama-3-8B), M in a list, with and without
the small-scale dequant (B200P_W_SMALLSCALE).  Each point is a CUDA graph of 8 x ncopy launches rotating over weight copies
larger than L2, device events around 3 replays; prints us per launch, GB/s of algorithmic weight bytes and the fraction of
the measured HBM peak (MEASURED_PEAKS.json).

    python scripts/qmm_layer_bench.py [Ms=1,8,64] [model=llama3-8b|qwen3-14b|llama3-8b-tp8|llama3-70b-8bit-tp8] [only=qkv,o,...]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlx_parallm_b200 import _lib, ops

L = _lib.lib()
_lib.require_device()
dt = torch.float16
g = torch.Generator(device="cuda").manual_seed(1)
Ms = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "1,8,64").split(",")]
model = sys.argv[2] if len(sys.argv) > 2 else "llama3-8b"
only = set(sys.argv[3].split(",")) if len(sys.argv) > 3 else None
SHAPES = {   # name, N, K, epilogue (0 store 16-bit, 1 store f32, 2 += f32, 3 silu*mul), bits
    "llama3-8b": [("qkv", 6144, 4096, 0, 4), ("o", 4096, 4096, 2, 4), ("gateup", 28672, 4096, 3, 4), ("down", 4096, 14336, 2, 4),
                  ("head", 128256, 4096, 1, 4)],
    "qwen3-14b": [("qkv", 7168, 5120, 0, 4), ("o", 5120, 5120, 2, 4), ("gateup", 34816, 5120, 3, 4), ("down", 5120, 17408, 2, 4)],
    # the fused projections through the fp32 accumulate epilogue (stream-K kernel) instead of their 16-bit epilogues
    "llama3-8b-acc": [("gateup", 28672, 4096, 2, 4), ("head", 128256, 4096, 2, 4), ("gateup-tp2", 14336, 4096, 2, 4),
                      ("gateup-tp4", 7168, 4096, 2, 4), ("gateup-tp8", 3584, 4096, 2, 4), ("head-tp8", 16032, 4096, 2, 4)],
    # accumulate-epilogue matrices of one GPU and of the TP=2 / TP=8 shards (the candidates of the mma.sync small-matrix GEMM)
    "small-acc": [("qkv", 6144, 4096, 2, 4), ("o", 4096, 4096, 2, 4), ("down", 4096, 14336, 2, 4),
                  ("qkv-tp2", 3072, 4096, 2, 4), ("o-tp2", 4096, 2048, 2, 4), ("down-tp2", 4096, 7168, 2, 4),
                  ("qkv-tp8", 768, 4096, 2, 4), ("o-tp8", 4096, 512, 2, 4), ("gateup-tp8", 3584, 4096, 2, 4),
                  ("down-tp8", 4096, 1792, 2, 4)],
    "llama3-8b-tp2": [("qkv", 3072, 4096, 0, 4), ("o", 4096, 2048, 2, 4), ("gateup", 14336, 4096, 3, 4), ("down", 4096, 7168, 2, 4),
                      ("head", 64128, 4096, 1, 4)],
    "llama3-8b-tp4": [("qkv", 1536, 4096, 0, 4), ("o", 4096, 1024, 2, 4), ("gateup", 7168, 4096, 3, 4), ("down", 4096, 3584, 2, 4),
                      ("head", 32064, 4096, 1, 4)],
    "llama3-8b-tp8": [("qkv", 768, 4096, 0, 4), ("o", 4096, 512, 2, 4), ("gateup", 3584, 4096, 3, 4), ("down", 4096, 1792, 2, 4),
                      ("head", 16032, 4096, 1, 4)],
    "llama3-70b-8bit-tp8": [("qkv", 1280, 8192, 0, 8), ("o", 8192, 1024, 2, 8), ("gateup", 7168, 8192, 3, 8), ("down", 8192, 3584, 2, 8)],
}[model]
peak = 6455.3
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
PATH = int(os.environ.get("QMM_BENCH_PATH", "2"))      # 2 tcgen05 tile kernels (default), 4 mma.sync small-matrix GEMM, 0 auto
_lib.check(L.b200p_qmm_set_path(PATH))
print(f"model {model}; path {PATH}; us per launch (GB/s, fraction of {peak:.0f} GB/s): 3-op dequant | small-scale dequant")
for name, N, K, epi, bits in SHAPES:
    if only and name not in only:
        continue
    wbytes = N * K * bits // 8 + 2 * N * (K // 64) * 2
    ncopy = max(2, min(8, int(4e8 // wbytes) + 1))
    mats = []
    for i in range(ncopy):
        w = torch.randint(-2**31, 2**31, (N, K * bits // 32), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        s = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
        b = (torch.rand((N, K // 64), generator=g, device="cuda") * 0.01).to(dt)
        mats.append((w, s, b))
    for M in Ms:
        x = torch.randn((M, K), device="cuda").to(dt)
        ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
        out = torch.zeros((M, N // 2 if epi == 3 else N), device="cuda", dtype=torch.float32 if epi in (1, 2) else dt)
        res = []
        for small in (False, True):
            reps = 8

            def body():
                for _ in range(reps):
                    for m in mats:
                        ops.qmm(x, *m, 64, bits, epi, out=out, workspace=ws, x_perm=(bits == 4), y_perm=(bits == 4 and epi in (0, 3)),
                                small_scale=small)
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
        print(f"{name:7s} N={N:6d} K={K:5d} b{bits} epi{epi} M={M:3d} : " +
              " | ".join(f"{r:6.1f} ({wbytes / r / 1e3:5.0f}, {wbytes / r / 1e3 / peak:.2f})" for r in res), flush=True)
    del mats
    torch.cuda.empty_cache()
_lib.check(L.b200p_qmm_set_path(0))
