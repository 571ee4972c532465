"""SURVEY 8c plan: the day ``import mlx.core`` works next to a reference checkout, these tests pin the oracle -- and through
it the CUDA path -- against the reference's own MLX run on the same synthetic checkpoint.  Skipped (with the reason) on
machines without MLX: parity stays "unpinned" there, as DESIGN.md section 2 states."""
import numpy as np
import pytest

from oracle import mlx_adapter

needs_mlx = pytest.mark.skipif(not mlx_adapter.available(),
                               reason="real mlx + the reference tree are not both present: oracle_kind = restatement")


def test_adapter_reports_what_it_is():
    kind = mlx_adapter.oracle_kind()
    assert kind == "restatement" or kind.startswith("mlx-cpu-")


@needs_mlx
@pytest.mark.parametrize("dtype", ["float32", "float16", "bfloat16"])
@pytest.mark.parametrize("bits", [4, 8])
def test_oracle_dequantize_and_quantize_are_bit_exact_against_mlx(dtype, bits):
    """Locks the rounding recipe the oracle (and with it b200p_dequantize, which is bit-exact against the oracle in every
    mode) assumes for mx.dequantize: exactly one of the oracle's modes must reproduce MLX bit for bit."""
    from oracle import mlx_ops as ref
    rng = np.random.default_rng(bits)
    N, K = 48, 512
    w = ref.round_to(rng.standard_normal((N, K)) * 0.05, dtype)
    q_o, s_o, b_o = ref.quantize(w, 64, bits, dtype)
    q_m, s_m, b_m = mlx_adapter.quantize(w.astype(np.float32 if dtype != "float16" else np.float16), 64, bits)
    assert np.array_equal(q_o.view(np.uint32), q_m.view(np.uint32))
    assert np.array_equal(np.asarray(s_o, np.float32), s_m) and np.array_equal(np.asarray(b_o, np.float32), b_m)
    got = mlx_adapter.dequantize(q_o, s_o, b_o, 64, bits)
    modes = {m: np.array_equal(np.asarray(ref.dequantize(q_o, s_o, b_o, 64, bits, dtype, mode=m), np.float32), got)
             for m in (ref.MODE_F32_MULADD, ref.MODE_FMA, ref.MODE_T_MULADD)}
    assert any(modes.values()), modes
    assert modes[ref.MODE_F32_MULADD], f"the default recipe is not MLX's: {modes} -- switch B200P_DQ default"


@needs_mlx
def test_oracle_logits_and_greedy_match_the_reference_run(tmp_path):
    from mlx_parallm_b200 import synth
    from oracle.model import OracleConfig, OracleModel, generate_step
    cfg = synth.get_config("tiny-llama")
    w = synth.random_weights(cfg, seed=1234, dtype="float16")
    synth.save_checkpoint(str(tmp_path), cfg, w)
    prompts = np.random.default_rng(7).integers(0, cfg["vocab_size"], (8, 16))
    ref_run = mlx_adapter.compare_all(str(tmp_path), prompts, steps=32)
    om = OracleModel(OracleConfig.from_hf(cfg, "float16"), synth.to_numpy_weights(w), numerics="reference")
    logits = om.forward(prompts, om.make_cache(8))
    assert np.abs(logits - ref_run["logits"]).max() / np.abs(ref_run["logits"]).max() < 1e-3
    toks = np.stack([y[:, 0] for _, (y, _) in zip(range(32), generate_step(prompts, om))], 1)
    assert (toks == ref_run["tokens"]).all(1).mean() >= 0.99


@needs_mlx
@pytest.mark.gpu
def test_cuda_logits_and_greedy_match_the_reference_run(tmp_path):
    import torch
    from mlx_parallm_b200 import synth
    from mlx_parallm_b200.utils import generate_step, load_model
    cfg = synth.get_config("small-llama")
    w = synth.random_weights(cfg, seed=1234, dtype="float16")
    synth.save_checkpoint(str(tmp_path), cfg, w)
    prompts = np.random.default_rng(7).integers(0, cfg["vocab_size"], (8, 16))
    ref_run = mlx_adapter.compare_all(str(tmp_path), prompts, steps=32)
    gm = load_model(tmp_path)
    got = gm(torch.from_numpy(prompts)).cpu().numpy()
    assert np.abs(got - ref_run["logits"]).max() / np.abs(ref_run["logits"]).max() < 1e-2
    toks = np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(32), generate_step(torch.from_numpy(prompts), gm))], 1)
    assert (toks == ref_run["tokens"]).all(1).mean() >= 0.99
