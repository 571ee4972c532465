"""LoRA at inference (row f4): b200p_lora_add against numpy, the adapted model against the oracle with the same adapters
(mlx_lm LoRALinear semantics), adapters loaded from the reference's on-disk format."""
import json

import numpy as np
import pytest
import torch

from helpers import make_pair, prompts_for, rel_err
from oracle import mlx_ops as ref
from oracle.model import generate_step as oracle_generate_step

pytestmark = pytest.mark.gpu
TD = {"float16": torch.float16, "bfloat16": torch.bfloat16}


def _pad_t(a, b, dtype):
    K, r = a.shape
    At = np.zeros((16, K), np.float32)
    At[:r] = a.T
    Bt = np.zeros((b.shape[1], 16), np.float32)
    Bt[:, :r] = b.T
    return torch.from_numpy(At).to(TD[dtype]).cuda(), torch.from_numpy(Bt).to(TD[dtype]).cuda()


@pytest.mark.parametrize("M,K,N,r,dtype,f32out", [(5, 256, 512, 16, "float16", True), (64, 4096, 4096, 8, "float16", True),
                                                  (33, 512, 1000, 16, "bfloat16", True), (70, 1024, 1024, 4, "float16", False)])
def test_lora_add_kernel(M, K, N, r, dtype, f32out):
    from mlx_parallm_b200 import ops
    rng = np.random.default_rng(M + K + r)
    x = ref.round_to(rng.standard_normal((M, K)), dtype)
    a = ref.round_to(rng.standard_normal((K, r)) / np.sqrt(K), dtype)
    b = ref.round_to(rng.standard_normal((r, N)) * 0.1, dtype)
    scale, col0, ld = 10.0, 128, N + 256
    y0 = rng.standard_normal((M, ld)).astype(np.float32)
    if not f32out:
        y0 = ref.round_to(y0, dtype)
    At, Bt = _pad_t(a, b, dtype)
    y = torch.from_numpy(y0.copy()).cuda() if f32out else torch.from_numpy(y0.copy()).to(TD[dtype]).cuda()
    ops.lora_add(torch.from_numpy(x).to(TD[dtype]).cuda(), At, Bt, y, col0, scale)
    t = ref.round_to(scale * (x @ a), dtype)                       # the kernel rounds scale * (x @ A) to the storage type
    upd = t @ b
    exp = y0.copy()
    exp[:, col0:col0 + N] += upd if f32out else ref.round_to(upd, dtype)
    if not f32out:
        exp = ref.round_to(exp, dtype)
    got = y.float().cpu().numpy()
    assert np.array_equal(got[:, :col0], y0[:, :col0]) and np.array_equal(got[:, col0 + N:], y0[:, col0 + N:])
    tol = {"float16": 2e-3, "bfloat16": 1.5e-2}[dtype]
    assert rel_err(got[:, col0:col0 + N] - y0[:, col0:col0 + N], exp[:, col0:col0 + N] - y0[:, col0:col0 + N]) < (tol if f32out else 5 * tol)


def _adapters(cfg, rank, rng, layers):
    D = cfg.get("head_dim") or cfg["hidden_size"] // cfg["num_attention_heads"]
    h = cfg["hidden_size"]
    out = {}
    for i in layers:
        for key, n in (("self_attn.q_proj", cfg["num_attention_heads"] * D), ("self_attn.v_proj", cfg["num_key_value_heads"] * D)):
            out[f"model.layers.{i}.{key}.lora_a"] = torch.from_numpy((rng.standard_normal((h, rank)) / np.sqrt(h)).astype(np.float32)).to(torch.float16)
            out[f"model.layers.{i}.{key}.lora_b"] = torch.from_numpy((rng.standard_normal((rank, n)) * 0.02).astype(np.float32)).to(torch.float16)
    return out


def test_model_with_lora_matches_oracle_and_loads_from_disk(tmp_path):
    from safetensors.torch import save_file
    from mlx_parallm_b200 import utils
    cfg, om, gm = make_pair("small-llama", dtype="float16")
    rng = np.random.default_rng(5)
    ad = _adapters(cfg, 8, rng, layers=[1])                # adapters on the last layer only (num_layers = 1)
    scale = 10.0
    for k in list(ad):
        if k.endswith("lora_a"):
            pre = k[:-len(".lora_a")]
            om.lora[pre] = (ad[k].float().numpy(), ad[pre + ".lora_b"].float().numpy(), scale)
    ids = prompts_for(cfg, 4, 30, seed=2)                    # 120 rows: the 16-bit prefill variant; decode: the fp32 variant
    base = gm(torch.from_numpy(ids), cache=gm.make_cache(4), last_only=True).cpu().numpy()
    # on-disk format of lora_init.py:140-149
    (tmp_path / "ad").mkdir()
    save_file({k: v.contiguous() for k, v in ad.items()}, str(tmp_path / "ad" / "adapters.safetensors"))
    json.dump({"fine_tune_type": "lora", "num_layers": 1,
               "lora_parameters": {"rank": 8, "scale": scale, "dropout": 0.0, "keys": ["self_attn.q_proj", "self_attn.v_proj"]}},
              open(tmp_path / "ad" / "adapter_config.json", "w"))
    utils.load_adapters(gm, tmp_path / "ad")
    exp = om.forward(ids, om.make_cache(4), last_only=True)
    got = gm(torch.from_numpy(ids), cache=gm.make_cache(4), last_only=True).cpu().numpy()
    assert rel_err(got, exp) < 1e-2
    assert rel_err(got, base) > 1e-3                        # the adapters do change the logits
    g = np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(6), utils.generate_step(torch.from_numpy(ids), gm))], 1)
    e = np.stack([y[:, 0] for _, (y, _) in zip(range(6), oracle_generate_step(ids, om))], 1)
    assert np.array_equal(g, e)
    gm.apply_lora({}, scale)                                # detach: back to the base model
    again = gm(torch.from_numpy(ids), cache=gm.make_cache(4), last_only=True).cpu().numpy()
    assert rel_err(again, base) < 1e-3                      # split-K partial sums arrive in any order: equal to fp32 rounding
