"""Model-level parity: logits within the north star's 1e-2 relative, greedy ids against the oracle and the
committed golden transcripts, and the reference-API behaviours of generate_step."""
import os

import numpy as np
import pytest
import torch

from helpers import greedy_report, make_pair, prompts_for, rel_err
from oracle.model import generate_step as oracle_generate_step

pytestmark = pytest.mark.gpu

LOGIT_TOL = 1e-2          # north star: per-step logits within 1e-2 relative (fp32 accumulate)


@pytest.mark.parametrize("name,dtype", [("tiny-llama", "float16"), ("tiny-qwen3", "bfloat16"),
                                        ("small-llama", "float16"), ("tiny-llama", "float32")])
def test_prefill_logits_all_positions(name, dtype):
    cfg, om, gm = make_pair(name, dtype=dtype)
    ids = prompts_for(cfg, 3, 11)
    exp = om.forward(ids, om.make_cache(3))
    got = gm(torch.from_numpy(ids), cache=None).cpu().numpy()
    assert got.shape == exp.shape == (3, 11, cfg["vocab_size"])
    err = rel_err(got, exp)
    print(f"{name} {dtype}: logits rel err {err:.2e}")
    assert err < LOGIT_TOL


@pytest.mark.parametrize("name,dtype,B,L", [("small-llama", "float16", 8, 40), ("tiny-qwen3", "bfloat16", 5, 33),
                                            ("small-llama", "float16", 2, 32), ("small-llama", "float16", 4, 32), ("small-llama", "float16", 3, 32),
                                            ("small-llama", "float16", 3, 43)])
def test_prefill_above_and_at_the_fp32_accumulator_limit(name, dtype, B, L):
    """B*L > ACC32_MAX_ROWS (128) rows take the 16-bit fused epilogues (STORE q|k|v, SILU_MUL gate|up), B*L <= 128 the fp32
    accumulators (64, 128 rows = whole tiles; 129 = first size above):
    same logits within tolerance, and the decode steps that follow (M = B rows) agree with the oracle."""
    cfg, om, gm = make_pair(name, dtype=dtype)
    ids = prompts_for(cfg, B, L, seed=3)
    exp = om.forward(ids, om.make_cache(B), last_only=True)
    got = gm(torch.from_numpy(ids), cache=gm.make_cache(B), last_only=True).cpu().numpy()
    err = rel_err(got, exp)
    print(f"{name} {dtype}: logits rel err {err:.2e}")
    assert err < LOGIT_TOL
    from mlx_parallm_b200.utils import generate_step
    g = np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(4), generate_step(torch.from_numpy(ids), gm))], 1)
    steps = [(y[:, 0], lg) for _, (y, lg) in zip(range(4), oracle_generate_step(ids, om))]
    e = np.stack([y for y, _ in steps], 1)
    frac, near, real = greedy_report(g, e, [lg for _, lg in steps], LOGIT_TOL)
    assert not real, real                         # a row may leave the oracle's transcript at a near-tie only


@pytest.mark.parametrize("name,dtype", [("tiny-llama", "float16"), ("small-llama", "float16"), ("tiny-qwen3", "bfloat16")])
def test_mirror_numerics_are_tight(name, dtype):
    """Against the oracle evaluated with the CUDA path's own rounding points the gap is far below 1e-2:
    what remains of the 1e-2 budget is the documented B3 deviation (16-bit KV / GEMM inputs), not kernel error."""
    cfg, om, gm = make_pair(name, dtype=dtype, numerics="mirror")
    ids = prompts_for(cfg, 2, 9)
    exp = om.forward(ids, om.make_cache(2), last_only=True)
    got = gm(torch.from_numpy(ids), cache=gm.make_cache(2), last_only=True).cpu().numpy()
    assert rel_err(got, exp) < (3e-3 if dtype == "float16" else 2e-2)


def run_greedy(gm, prompts, steps):
    from mlx_parallm_b200.utils import generate_step
    toks = []
    for _, (y, p) in zip(range(steps), generate_step(torch.from_numpy(prompts), gm)):
        assert y.shape == (prompts.shape[0], 1) and p.shape == (prompts.shape[0], 1)
        toks.append(y[:, 0].cpu().numpy())
    return np.stack(toks, 1)


def sequences_identical(got, exp):
    return (got == exp).all(1)


@pytest.mark.parametrize("fixture,name,dtype", [("tiny_llama_f16_greedy", "tiny-llama", "float16"),
                                                ("tiny_qwen3_bf16_greedy", "tiny-qwen3", "bfloat16"),
                                                ("small_llama_f16_greedy", "small-llama", "float16"),
                                                ("tiny_f32_greedy", "tiny", "float32")])
def test_greedy_tokens_match_golden_transcript(golden_dir, fixture, name, dtype):
    g = np.load(os.path.join(golden_dir, fixture + ".npz"))
    cfg, om, gm = make_pair(name, dtype=dtype)
    got = run_greedy(gm, g["prompts"], 32)
    same = sequences_identical(got, g["tokens"])
    if not same.all():
        # EVERY row that leaves the golden transcript must do so at a near-tie of the oracle (logit margin of the engine's
        # token below the 1e-2 relative logits tolerance); anything else is a kernel bug
        logits = [lg for _, (_, lg) in zip(range(32), oracle_generate_step(g["prompts"], om))]
        frac, near, real = greedy_report(got, g["tokens"], logits, LOGIT_TOL)
        assert not real, real
    if dtype != "bfloat16":            # bf16 (8 mantissa bits) is held to the logits bar and the near-tie rule, see DESIGN 2
        assert same.mean() >= 0.99, same


def test_decode_logits_per_step_and_graph_equals_eager():
    from mlx_parallm_b200 import utils
    cfg, om, gm = make_pair("small-llama", dtype="float16")
    p = prompts_for(cfg, 4, 20)
    # teacher-forced: feed the ORACLE's tokens to both so every step compares like with like
    cache_o = om.make_cache(4)
    cache_g = gm.make_cache(4)
    lo = om.forward(p, cache_o, last_only=True)[:, -1]
    lg = gm(torch.from_numpy(p), cache=cache_g, last_only=True)[:, -1].cpu().numpy()
    assert rel_err(lg, lo) < LOGIT_TOL
    step = utils._DecodeStep(gm, cache_g, 4, greedy=True, use_graph=True)
    step_e = None
    for t in range(6):
        y = lo.argmax(-1)
        lo = om.forward(y[:, None], cache_o, last_only=True)[:, -1]
        step.tok.copy_(torch.from_numpy(y.astype(np.int32)))
        step.run()
        assert rel_err(step.logits.cpu().numpy(), lo) < LOGIT_TOL, t
        assert step.tok.cpu().numpy().tolist() == step.logits.argmax(-1).cpu().numpy().tolist()
    assert step.graph is not None and cache_g[0].offsets == [26] * 4
    assert cache_g[0].lengths("cuda").tolist() == [26] * 4


def test_generate_step_api_behaviours():
    from mlx_parallm_b200.utils import generate_step
    cfg, om, gm = make_pair("tiny-llama", dtype="float16")
    p = prompts_for(cfg, 2, 6)
    with pytest.raises(NotImplementedError):
        next(generate_step(p, gm, repetition_penalty=1.1))
    # cache growth across the 256-token step boundary keeps the sequence identical to a big-step cache
    c_small = gm.make_cache(2)
    for c in c_small:
        c.step = 8
    a = [y for _, (y, _) in zip(range(12), generate_step(p, gm, cache=c_small))]
    b = [y for _, (y, _) in zip(range(12), generate_step(p, gm, cache=gm.make_cache(2)))]
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    assert c_small[0].keys.shape[2] >= 18 and c_small[0].keys.shape[2] % 8 == 0
    # logit_bias forces a token (utils.py:346-349)
    y, _ = next(generate_step(p, gm, logit_bias={7: 1e4}))
    assert y.view(-1).tolist() == [7, 7]
    # sampling path: tokens valid, reproducible under a seed
    from mlx_parallm_b200 import sample_utils
    sample_utils.seed(3)
    s1 = [y for _, (y, _) in zip(range(5), generate_step(p, gm, temp=0.8, top_p=0.9))]
    sample_utils.seed(3)
    s2 = [y for _, (y, _) in zip(range(5), generate_step(p, gm, temp=0.8, top_p=0.9))]
    assert all(torch.equal(x, y) for x, y in zip(s1, s2))
    assert all(0 <= int(v) < cfg["vocab_size"] for t in s1 for v in t.view(-1))


def test_load_model_from_mlx_checkpoint(tmp_path):
    from mlx_parallm_b200 import synth, utils
    cfg = synth.get_config("tiny-qwen3")
    w = synth.random_weights(cfg, seed=9, dtype="float16")
    w["model.layers.0.self_attn.rotary_emb.inv_freq"] = torch.zeros(4)       # sanitize / unknown-key filter
    synth.save_checkpoint(str(tmp_path), cfg, w)
    m = utils.load_model(tmp_path)
    _, om, gm = make_pair("tiny-qwen3", dtype="float16", seed=9)
    ids = prompts_for(cfg, 2, 5)
    assert torch.equal(m(torch.from_numpy(ids)), gm(torch.from_numpy(ids)))
    assert m.head_dim == 64 and m.n_kv_heads == 2 and len(m.layers) == 3 and m.model_type == "qwen3"


def test_phi3_checkpoint_with_fused_projections(tmp_path):
    """Phi-3 (reference models/phi3.py): the checkpoint holds qkv_proj / gate_up_proj; logits and greedy ids against the
    oracle's restatement of the same graph, loaded through load_model from the fused keys."""
    from mlx_parallm_b200 import synth, utils
    from oracle.model import OracleConfig, OracleModel, split_phi3_weights
    cfg = synth.get_config("small-phi3")
    ck = synth.to_phi3_checkpoint(synth.random_weights(cfg, seed=21, dtype="float16"))
    synth.save_checkpoint(str(tmp_path), cfg, ck)
    gm = utils.load_model(tmp_path)
    assert gm.model_type == "phi3" and gm.head_dim == 128 and gm.n_kv_heads == 2 and gm.lm_head is not None
    om = OracleModel(OracleConfig.from_hf(cfg, "float16"), split_phi3_weights(synth.to_numpy_weights(ck), cfg))
    ids = prompts_for(cfg, 5, 19)
    exp = om.forward(ids, om.make_cache(5), last_only=True)
    got = gm(torch.from_numpy(ids), cache=gm.make_cache(5), last_only=True).cpu().numpy()
    assert rel_err(got, exp) < LOGIT_TOL
    g = run_greedy(gm, ids, 16)
    e = np.stack([y[:, 0] for _, (y, _) in zip(range(16), oracle_generate_step(ids, om))], 1)
    assert sequences_identical(g, e).all()


def test_convert_quantizes_a_dense_checkpoint_like_the_oracle(tmp_path):
    """utils.convert(quantize=True): HF dense safetensors -> MLX 4-bit checkpoint whose tensors equal the oracle's
    mx.quantize restatement bit for bit, and which load_model + generate_step run."""
    from safetensors.torch import save_file
    import json
    from mlx_parallm_b200 import synth, utils
    from oracle import mlx_ops as ref
    cfg = synth.get_config("tiny-llama")
    dense_cfg = {k: v for k, v in cfg.items() if k != "quantization"}
    rng = np.random.default_rng(21)
    dense = {}
    for name, (n, k) in synth.linear_shapes(cfg).items():
        dense[name + ".weight"] = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32) / np.sqrt(k)).to(torch.float16)
    for name in synth.norm_names(cfg):
        dense[name] = torch.ones((cfg["hidden_size"],), dtype=torch.float16)
    src = tmp_path / "hf"
    src.mkdir()
    save_file({k: v.contiguous() for k, v in dense.items()}, str(src / "model.safetensors"))
    json.dump(dense_cfg, open(src / "config.json", "w"))
    out = utils.convert(str(src), str(tmp_path / "mlx"), quantize=True, q_group_size=64, q_bits=4)
    from safetensors import safe_open
    with safe_open(str(out / "model.safetensors"), framework="pt") as f:
        assert f.metadata() == {"format": "mlx"}
        got = {k: f.get_tensor(k) for k in f.keys()}
    assert json.load(open(out / "config.json"))["quantization"] == {"group_size": 64, "bits": 4}
    name = "model.layers.1.mlp.down_proj"
    eq, es, eb = ref.quantize(dense[name + ".weight"].float().numpy(), 64, 4, "float16")
    assert np.array_equal(got[name + ".weight"].view(torch.int32).numpy().view(np.uint32), eq)
    assert np.array_equal(got[name + ".scales"].float().numpy(), es) and np.array_equal(got[name + ".biases"].float().numpy(), eb)
    gm = utils.load_model(out)
    y, _ = next(utils.generate_step(torch.zeros((2, 5), dtype=torch.int64), gm))
    assert y.shape == (2, 1)
