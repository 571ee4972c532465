"""Regenerates tests/golden/*.npz from the oracle (run from the repo root:  python tests/golden/make_golden.py).

The reference cannot be imported here (mlx is not installable), so these fixtures pin the ORACLE against
regressions and give the CUDA path fixed targets; they are not outputs of real MLX (parity unpinned).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import make_pair, prompts_for  # noqa: E402
from oracle import mlx_ops as ops  # noqa: E402
from oracle.model import generate_step  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def transcript(name, dtype, B, Lp, steps, **ov):
    cfg, om, _ = make_pair(name, dtype=dtype, device=None, **ov)
    prompts = prompts_for(cfg, B, Lp)
    toks, first_logits = [], None
    for i, (y, lg) in zip(range(steps), generate_step(prompts, om)):
        toks.append(y[:, 0])
        if i == 0:
            first_logits = lg
    fl = first_logits.astype(np.float32)
    stride = max(1, fl.shape[1] // 2048)          # keep fixtures small: every stride-th column + the argmax
    return dict(prompts=prompts, tokens=np.stack(toks, 1), first_logits=fl[:, ::stride],
                logit_stride=np.int64(stride), first_max=fl.max(-1), first_absmax=np.abs(fl).max(-1))


def main():
    # 1. greedy transcripts: config 1 "tiny" (f32, V=151936, B=8, 16-token prompts, 32 tokens)
    np.savez_compressed(os.path.join(OUT, "tiny_f32_greedy.npz"), **transcript("tiny", "float32", 8, 16, 32))
    np.savez_compressed(os.path.join(OUT, "tiny_llama_f16_greedy.npz"), **transcript("tiny-llama", "float16", 4, 12, 32))
    np.savez_compressed(os.path.join(OUT, "tiny_qwen3_bf16_greedy.npz"), **transcript("tiny-qwen3", "bfloat16", 4, 12, 32))
    np.savez_compressed(os.path.join(OUT, "small_llama_f16_greedy.npz"), **transcript("small-llama", "float16", 4, 20, 32))
    np.savez_compressed(os.path.join(OUT, "small_phi3_f16_greedy.npz"), **transcript("small-phi3", "float16", 4, 20, 32))
    # 2. dequantize bit patterns: every code x random scales, 3 dtypes x 2 widths x 3 modes
    rng = np.random.default_rng(99)
    dq = {}
    for bits in (4, 8):
        n_codes = 1 << bits
        K = 64 * ((n_codes + 63) // 64)
        N = 32
        q = (np.arange(N * K).reshape(N, K) % n_codes).astype(np.uint32)
        wq = ops.pack_codes(q, bits)
        for dt in ops.DTYPES:
            s = ops.round_to(rng.uniform(-0.05, 0.05, (N, K // 64)).astype(np.float32), dt)
            b = ops.round_to(rng.uniform(-0.5, 0.5, (N, K // 64)).astype(np.float32), dt)
            dq[f"w_{bits}"] = wq
            dq[f"s_{bits}_{dt}"] = s
            dq[f"b_{bits}_{dt}"] = b
            for mode in (0, 1, 2):
                out = ops.dequantize(wq, s, b, 64, bits, dt, mode)
                dq[f"out_{bits}_{dt}_{mode}"] = ops.storage_bits(out, dt)
    np.savez_compressed(os.path.join(OUT, "dequantize_bits.npz"), **dq)
    # 3. quantize: codes + scale / bias bit patterns for dense inputs covering both sign branches, exact zeros, all-zero groups
    rngq = np.random.default_rng(77)
    qz = {}
    wd = (rngq.standard_normal((24, 256)) * 0.05).astype(np.float32)
    wd[0] = 0.0
    wd[1, :64] = -np.abs(wd[1, :64])
    wd[2, :64] = np.abs(wd[2, :64])
    wd[3, ::2] = 0.0
    for dt in ops.DTYPES:
        w = ops.round_to(wd, dt)
        qz[f"w_{dt}"] = ops.storage_bits(w, dt)
        for bits in (4, 8):
            for gs in (32, 64, 128):
                wq, sc, bi = ops.quantize(w, gs, bits, dt)
                qz[f"q_{dt}_{bits}_{gs}"] = wq
                qz[f"s_{dt}_{bits}_{gs}"] = ops.storage_bits(sc, dt)
                qz[f"b_{dt}_{bits}_{gs}"] = ops.storage_bits(bi, dt)
    np.savez_compressed(os.path.join(OUT, "quantize_bits.npz"), **qz)
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
