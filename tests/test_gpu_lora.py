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


def test_lora_on_a_tensor_parallel_shard_keeps_its_output_columns():
    """q_proj / v_proj are column-parallel: a rank applies lora_a whole and the columns of lora_b that belong to its rows of
    the projection (the range tp.shard_weights keeps), so its q|k|v block equals the same columns of the unsharded model."""
    from mlx_parallm_b200 import ops, synth, tp
    from mlx_parallm_b200.utils import _get_classes
    cfg = synth.get_config("small-llama", num_key_value_heads=2)      # the stock config has one kv head: not divisible
    w = synth.random_weights(cfg, seed=1234, dtype="float16")
    Model, Args = _get_classes(cfg)
    full = Model(Args.from_dict(cfg)).load_weights(w, device="cuda")
    rng = np.random.default_rng(11)
    ad = _adapters(cfg, 8, rng, layers=[0, 1])
    full.apply_lora(ad, 10.0)
    x = (torch.randn((6, cfg["hidden_size"]), generator=torch.Generator().manual_seed(3)) * 0.5).to(torch.float16).cuda()
    world = 2
    for rank in range(world):
        part = Model(Args.from_dict(cfg)).load_weights(tp.shard_weights(w, cfg, rank, world), device="cuda")
        part.set_tensor_parallel(rank, world, None)
        part.apply_lora(ad, 10.0)                                   # full-width adapters, cut inside
        for li in (0, 1):
            af, ap = full.model.layers[li].self_attn, part.model.layers[li].self_attn
            for name in ("q_proj", "v_proj"):
                n_loc = getattr(ap, name).weight.shape[0]
                Atf, Btf, sf = af.lora[name]
                Atp, Btp, sp = ap.lora[name]
                assert torch.equal(Atf, Atp) and sf == sp
                assert torch.equal(Btp, Btf[rank * n_loc:(rank + 1) * n_loc])
                yf = torch.zeros((6, Btf.shape[0]), dtype=torch.float32, device="cuda")
                yp = torch.zeros((6, n_loc), dtype=torch.float32, device="cuda")
                ops.lora_add(x, Atf, Btf, yf, 0, sf)
                ops.lora_add(x, Atp, Btp, yp, 0, sp)
                assert torch.equal(yp, yf[:, rank * n_loc:(rank + 1) * n_loc])


def test_lora_hot_swap_recaptures_the_decode_graph():
    """weight_updater.py:66-73: adapters are swapped while the server keeps generating.  The captured decode graph holds the
    old adapter pointers; after a swap the next generation must run on the new set (graph re-captured), and detaching must
    give the base model back."""
    from mlx_parallm_b200 import utils
    cfg, om, gm = make_pair("small-llama", dtype="float16")
    ids = torch.from_numpy(prompts_for(cfg, 2, 12, seed=4))

    def run():
        return np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(8), utils.generate_step(ids, gm))], 1)

    def logits():
        return gm(ids, cache=gm.make_cache(2), last_only=True).cpu().numpy()

    base_t, base_l = run(), logits()
    rng = np.random.default_rng(7)
    ad1, ad2 = _adapters(cfg, 8, rng, layers=[0, 1]), _adapters(cfg, 16, rng, layers=[1])
    gm.apply_lora(ad1, 10.0)
    l1, t1 = logits(), run()
    gm.apply_lora(ad2, 10.0)
    l2, t2 = logits(), run()
    assert rel_err(l1, base_l) > 1e-3 and rel_err(l2, base_l) > 1e-3 and rel_err(l1, l2) > 1e-3
    fresh = make_pair("small-llama", dtype="float16")[2].apply_lora(ad2, 10.0)
    tf = np.stack([y[:, 0].cpu().numpy() for _, (y, _) in zip(range(8), utils.generate_step(ids, fresh))], 1)
    lf = fresh(ids, cache=fresh.make_cache(2), last_only=True).cpu().numpy()
    assert rel_err(l2, lf) < 1e-4 and (t2[:, 0] == tf[:, 0]).all()          # same set on a fresh model (fp32 order aside)
    gm.apply_lora({}, 10.0)
    assert rel_err(logits(), base_l) < 1e-3
    del t1, base_t
