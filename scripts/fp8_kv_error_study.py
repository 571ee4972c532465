"""What would an fp8 (E4M3) KV cache cost against the north star's 1e-2 relative logits tolerance?  CPU study with the oracle:
the same model and prompts with the K/V cache stored as float32 (the reference, base.py:111-112), float16 (this build) and
E4M3 without / with one fp32 scale per (token, kv head) -- logits of the last position and of 8 greedy decode steps.

    python scripts/fp8_kv_error_study.py        (about a minute on 8 cores; no GPU)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

from mlx_parallm_b200 import synth
from oracle import mlx_ops as ops
from oracle.model import OracleConfig, OracleModel, OraclePagedKVCache


class ScaledE4M3Cache(OraclePagedKVCache):
    """E4M3 codes with one scale per appended (row, kv head, token) vector: x -> round(x / s) * s, s = max|x| / 448."""

    def update_and_fetch(self, k, v):
        def q(x):
            s = np.maximum(np.abs(x).max(-1, keepdims=True), 1e-30) / 448.0
            return ops.e4m3_round(x / s) * s
        return super().update_and_fetch(q(np.asarray(k, np.float32)), q(np.asarray(v, np.float32)))


def run(model, prompts, kind, steps=8, forced=None):
    """forced: the reference run's greedy ids per step (teacher forcing keeps the steps comparable after an id flips)"""
    B = prompts.shape[0]
    c = model.cfg
    if kind == "e4m3-scaled":
        cache = [ScaledE4M3Cache(c.head_dim, c.num_key_value_heads, B, "float32") for _ in range(c.num_hidden_layers)]
    else:
        cache = [OraclePagedKVCache(c.head_dim, c.num_key_value_heads, B, kind) for _ in range(c.num_hidden_layers)]
    out = []
    lg = model.forward(prompts, cache, last_only=True)[:, -1]
    out.append(lg)
    for t in range(steps):
        y = lg.argmax(-1) if forced is None else forced[t]
        lg = model.forward(y[:, None], cache, last_only=True)[:, -1]
        out.append(lg)
    return out


def main():
    cfg = synth.get_config("small-llama", num_hidden_layers=8, hidden_size=1024, intermediate_size=2048, num_attention_heads=8,
                           num_key_value_heads=2, vocab_size=2048)
    w = synth.random_weights(cfg, seed=1234, dtype="float16")
    model = OracleModel(OracleConfig.from_hf(cfg, "float16"), synth.to_numpy_weights(w))
    prompts = np.random.default_rng(7).integers(0, cfg["vocab_size"], (4, 192)).astype(np.int64)
    ref = run(model, prompts, "float32")
    print(f"model: {cfg['num_hidden_layers']} layers, hidden {cfg['hidden_size']}, GQA {cfg['num_attention_heads']}/{cfg['num_key_value_heads']}, "
          f"D=128, prompts 4 x 192, 8 greedy steps; relative = max|a-b| / max|b| per step, worst step reported")
    for kind in ("float16", "bfloat16", "float8_e4m3", "e4m3-scaled"):
        got = run(model, prompts, kind, forced=[r.argmax(-1) for r in ref[:-1]])
        errs = [float(np.abs(g - r).max() / np.abs(r).max()) for g, r in zip(got, ref)]
        same = np.mean([np.array_equal(g.argmax(-1), r.argmax(-1)) for g, r in zip(got, ref)])
        print(f"KV stored as {kind:12s}: logits rel err vs float32 KV: prefill {errs[0]:.2e}, worst decode step {max(errs[1:]):.2e}; "
              f"steps (teacher-forced) whose argmax equals the float32-KV argmax on all rows {same:.2f}")


if __name__ == "__main__":
    main()
