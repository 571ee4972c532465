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
         ("small_llama_f16_greedy", "small-llama", "float16")]


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
