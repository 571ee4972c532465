"""The oracle against the committed fixtures (tests/golden/make_golden.py): regression pin for the
restatement, and for the sizes the reference's own fixture uses (BASELINE config 1 'tiny')."""
import os

import numpy as np
import pytest

from oracle import mlx_ops as ops
from oracle.model import generate_step
from helpers import make_pair

CASES = [("tiny_llama_f16_greedy", "tiny-llama", "float16"),
         ("tiny_qwen3_bf16_greedy", "tiny-qwen3", "bfloat16"),
         ("small_llama_f16_greedy", "small-llama", "float16"),
         ("small_phi3_f16_greedy", "small-phi3", "float16")]


@pytest.mark.parametrize("fixture,name,dtype", CASES)
def test_oracle_greedy_transcript_matches_golden(golden_dir, fixture, name, dtype):
    g = np.load(os.path.join(golden_dir, fixture + ".npz"))
    cfg, om, _ = make_pair(name, dtype=dtype, device=None)
    n = 8                                        # first 8 of the 32 stored steps keeps the CPU suite quick
    toks = [y[:, 0] for _, (y, _) in zip(range(n), generate_step(g["prompts"], om))]
    assert np.array_equal(np.stack(toks, 1), g["tokens"][:, :n])


def test_oracle_tiny_config1_first_step_matches_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "tiny_f32_greedy.npz"))
    cfg, om, _ = make_pair("tiny", dtype="float32", device=None)
    y, lg = next(generate_step(g["prompts"], om))
    assert np.array_equal(y[:, 0], g["tokens"][:, 0])
    assert np.allclose(lg[:, ::int(g["logit_stride"])], g["first_logits"], rtol=0, atol=1e-4)


def test_oracle_dequantize_bits_match_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "dequantize_bits.npz"))
    for bits in (4, 8):
        for dt in ops.DTYPES:
            for mode in (0, 1, 2):
                out = ops.dequantize(g[f"w_{bits}"], g[f"s_{bits}_{dt}"], g[f"b_{bits}_{dt}"], 64, bits, dt, mode)
                assert np.array_equal(ops.storage_bits(out, dt), g[f"out_{bits}_{dt}_{mode}"]), (bits, dt, mode)


def _from_bits(bits, dt):
    if dt == "float32":
        return bits.view(np.float32)
    if dt == "float16":
        return bits.view(np.float16).astype(np.float32)
    return ops.bf16_from_bits(bits)


def test_oracle_quantize_bits_match_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "quantize_bits.npz"))
    for dt in ops.DTYPES:
        w = _from_bits(g[f"w_{dt}"], dt)
        for bits in (4, 8):
            for gs in (32, 64, 128):
                wq, sc, bi = ops.quantize(w, gs, bits, dt)
                assert np.array_equal(wq, g[f"q_{dt}_{bits}_{gs}"]), (dt, bits, gs)
                assert np.array_equal(ops.storage_bits(sc, dt), g[f"s_{dt}_{bits}_{gs}"])
                assert np.array_equal(ops.storage_bits(bi, dt), g[f"b_{dt}_{bits}_{gs}"])
                # properties of A.1: the scale is snapped so that zero sits on the code grid (up to storage rounding of
                # q*s + b), and codes stay in range
                d = ops.dequantize(wq, sc, bi, gs, bits, dt)
                assert np.abs(d[w == 0]).max() <= np.abs(w).max() * 2.0 ** -6
                assert ops.unpack_codes(wq, bits).max() <= (1 << bits) - 1
