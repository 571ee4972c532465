"""Host-side logic that needs no GPU: args, caches, byte model, synthetic checkpoints, detokenizer."""
import json
import os

import numpy as np
import pytest
import torch

from mlx_parallm_b200 import synth
from mlx_parallm_b200.models import llama, qwen3
from mlx_parallm_b200.models.base import BatchedKVCache, PagedKVCache
from mlx_parallm_b200.tokenizer_utils import NaiveStreamingDetokenizer, TokenizerWrapper


def test_model_args_from_dict_drops_unknown_keys_and_defaults():
    cfg = synth.get_config("mistral-7b")                 # carries sliding_window, which ModelArgs lacks
    a = llama.ModelArgs.from_dict(cfg)
    assert a.num_key_value_heads == 8 and not hasattr(a, "sliding_window")
    cfg2 = {k: v for k, v in cfg.items() if k not in ("num_key_value_heads", "tie_word_embeddings")}
    a2 = llama.ModelArgs.from_dict(cfg2)
    assert a2.num_key_value_heads == 32 and a2.tie_word_embeddings is True     # llama.py:30,33-34
    with pytest.raises(ValueError):
        llama.ModelArgs.from_dict(dict(cfg, rope_scaling={"type": "linear"}))   # no 'factor'
    with pytest.raises(ValueError):
        llama.ModelArgs.from_dict(dict(cfg, rope_scaling={"type": "yarn", "factor": 2}))
    q = qwen3.ModelArgs.from_dict(synth.get_config("qwen3-14b"))
    assert q.head_dim == 128


def test_phi3_fused_checkpoint_keys_split_into_the_projection_blocks():
    """models/phi3.py: qkv_proj rows [q | k | v], gate_up_proj rows [gate | up] (reference phi3.py:92-96, 123-125); the
    CUDA model's splitter and the oracle's agree with the per-projection tensors the checkpoint was fused from."""
    from mlx_parallm_b200.models import phi3
    from mlx_parallm_b200.utils import _get_classes
    from oracle.model import OracleConfig, split_phi3_weights
    cfg = synth.get_config("small-phi3")
    Model, Args = _get_classes(cfg)
    assert Model is phi3.Model
    a = Args.from_dict(dict(cfg, sliding_window=2047))
    assert a.tie_word_embeddings is False and a.num_key_value_heads == 2
    w = synth.random_weights(cfg, seed=5, dtype="float16")
    ck = synth.to_phi3_checkpoint(w)
    assert "model.layers.0.self_attn.qkv_proj.weight" in ck and "model.layers.1.mlp.gate_up_proj.scales" in ck
    assert not any(".q_proj." in k or ".up_proj." in k for k in ck)
    assert ck["model.layers.0.self_attn.qkv_proj.weight"].shape == ((4 + 2 * 2) * 128, 512 // 8)
    back = Model(a).split_fused(ck)
    assert set(back) == set(w) and all(torch.equal(back[k], w[k]) for k in w)
    ob = split_phi3_weights(synth.to_numpy_weights(ck), cfg)
    wn = synth.to_numpy_weights(w)
    assert set(ob) == set(wn) and all(np.array_equal(ob[k], wn[k]) for k in wn)
    oc = OracleConfig.from_hf(cfg, "float16")
    assert oc.model_type == "llama" and oc.tie_word_embeddings is False and oc.head_dim == 128
    with pytest.raises(NotImplementedError):                               # su_rope.py is not built
        Args.from_dict(dict(cfg, rope_scaling={"type": "su", "long_factor": [1.0] * 64, "short_factor": [1.0] * 64}))
    with pytest.raises(ValueError):
        Args.from_dict(dict(cfg, rope_scaling={"type": "su"}))
    bad = dict(ck)
    bad["model.layers.0.self_attn.qkv_proj.weight"] = bad["model.layers.0.self_attn.qkv_proj.weight"][:-8]
    with pytest.raises(ValueError):
        Model(a).split_fused(bad)


def test_byte_model_matches_baseline_md():
    gb = lambda n: synth.weight_bytes_per_step(synth.get_config(n)) / 1e9
    assert gb("mistral-7b") == pytest.approx(4.000, abs=2e-3)
    assert gb("llama-3-8b") == pytest.approx(4.222, abs=2e-3)
    assert gb("qwen3-14b") == pytest.approx(7.870, abs=2e-3)
    assert gb("llama-3-70b") == pytest.approx(73.848, abs=2e-3)
    assert synth.kv_bytes_per_token(synth.get_config("llama-3-8b")) == 131072
    assert synth.kv_bytes_per_token(synth.get_config("qwen3-14b")) == 163840
    assert synth.kv_bytes_per_token(synth.get_config("llama-3-70b")) == 327680


@pytest.mark.parametrize("klass", [BatchedKVCache, PagedKVCache])
def test_cache_surface_on_cpu_tensors(klass):
    c = klass(8, 2, batch_size=3)
    assert c.keys is None and c.offset == 0 and c.step == 256 and c.offsets == [0, 0, 0]
    c.step = 4
    k = torch.arange(3 * 2 * 3 * 8, dtype=torch.float32).reshape(3, 2, 3, 8)
    kf, vf = c.update_and_fetch(k, -k)
    assert kf.shape == (3, 2, 3, 8) and c.keys.shape[2] == 4 and c.offsets == [3, 3, 3] and c.offset == 3
    assert torch.equal(kf, k) and torch.equal(vf, -k)
    kf, _ = c.update_and_fetch(k[:, :, :2], k[:, :, :2])
    assert kf.shape[2] == 5 and c.keys.shape[2] == 8 and torch.equal(kf[:, :, :3], k)
    assert c.lengths("cpu").tolist() == [5, 5, 5]
    buf = c.keys
    c.reset()
    assert c.offsets == [0, 0, 0] and c.keys is buf and c.lengths("cpu").tolist() == [0, 0, 0]
    c.reset(2)
    assert c.keys is None and c.batch_size == 2 and c.offsets == [0, 0]


def test_synthetic_checkpoint_roundtrip(tmp_path):
    cfg = synth.get_config("tiny-llama")
    w = synth.random_weights(cfg, seed=5, dtype="bfloat16")
    w2 = synth.random_weights(cfg, seed=5, dtype="bfloat16", only=lambda n: "layers.1." in n)
    assert all(torch.equal(w[k], v) for k, v in w2.items()) and len(w2) == 7 * 3 + 2
    synth.save_checkpoint(str(tmp_path), cfg, w)
    from safetensors import safe_open
    with safe_open(os.path.join(tmp_path, "model.safetensors"), framework="pt") as f:
        assert f.metadata() == {"format": "mlx"}
        t = f.get_tensor("model.layers.0.mlp.down_proj.weight")
        assert t.dtype == torch.uint32 and t.shape == (128, 256 * 4 // 32)
        assert f.get_tensor("model.layers.0.mlp.down_proj.scales").shape == (128, 4)
    assert json.load(open(os.path.join(tmp_path, "config.json")))["quantization"] == {"group_size": 64, "bits": 4}


class _FakeTok:
    eos_token_id = 0
    eos_token = "<eos>"
    pad_token = None

    def decode(self, ids, **kw):
        return "".join(chr(96 + i) if i else "<eos>" for i in ids)


def test_streaming_detokenizer_and_wrapper_passthrough():
    tw = TokenizerWrapper(_FakeTok())
    assert tw.eos_token_id == 0 and tw._tokenizer.eos_token == "<eos>" and tw.eos_token_ids == {0}
    d = tw.detokenizer
    d.reset()
    segs = []
    for t in (1, 2, 3):
        d.add_token(t)
        segs.append(d.last_segment)
    d.finalize()
    segs.append(d.last_segment)
    assert "".join(segs) == "abc" == d.text and d.tokens == [1, 2, 3]
    tw.pad_token = "<eos>"                                 # setattr passes through (utils.py:513-515)
    assert tw._tokenizer.pad_token == "<eos>"


def test_compat_install_aliases_the_reference_module_names():
    """the names the reference's server / demo import resolve to this engine."""
    import sys
    import mlx_parallm_b200.compat as compat
    saved = {k: v for k, v in sys.modules.items() if k == "mlx_parallm" or k.startswith("mlx_parallm.")}
    try:
        compat.install()
        from mlx_parallm.utils import load, generate, stream_generate, batch_generate, generate_step, \
            batch_stream_generate_text, batch_generate_text                                   # noqa: F401  (server/main.py:22-28)
        from mlx_parallm.sample_utils import top_p_sampling                                    # noqa: F401  (server/main.py:29)
        from mlx_parallm.models.base import BatchedKVCache, PagedKVCache                       # noqa: F401
        import mlx_parallm.models.llama as llama
        import mlx_parallm_b200.utils as ours
        assert generate_step is ours.generate_step and llama.Model.__module__ == "mlx_parallm_b200.models.llama"
    finally:
        for k in [k for k in sys.modules if k == "mlx_parallm" or k.startswith("mlx_parallm.")]:
            del sys.modules[k]
        sys.modules.update(saved)


def test_encode_and_chat_template_lru_caches():
    """utils.py:137-197: one tokenizer call per distinct (tokenizer, text) / (tokenizer, messages, flag); LRU eviction."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from fake_tokenizer import FakeTokenizer
    from mlx_parallm_b200 import utils
    tw = TokenizerWrapper(FakeTokenizer(64))
    assert utils.encode_cached(tw, "w3 w4") == [3, 4] and utils.encode_cached(tw, "w3 w4") == [3, 4]
    assert tw._tokenizer.encode_calls == 1
    msgs = [{"role": "user", "content": "w5 w6", "extra": "ignored by the key"}]
    a = utils.apply_chat_template_cached(tw, msgs)
    b = utils.apply_chat_template_cached(tw, [{"content": "w5 w6", "role": "user"}])          # same (role, content)
    c = utils.apply_chat_template_cached(tw, msgs, add_generation_prompt=False)
    assert a == b == "w1 w5 w6 w2" and c == "w1 w5 w6" and tw._tokenizer.chat_template_calls == 2
    tw2 = TokenizerWrapper(FakeTokenizer(64))                                                 # another tokenizer: another key
    utils.apply_chat_template_cached(tw2, msgs)
    assert tw2._tokenizer.chat_template_calls == 1
    lru = utils._LRUCache(max_size=2)
    lru.set("a", 1); lru.set("b", 2); lru.get("a"); lru.set("c", 3)
    assert lru.get("b") is None and lru.get("a") == 1 and lru.get("c") == 3


def test_compat_provides_every_engine_name_the_reference_server_imports():
    """server/main.py:22-33 and rl_training import these from mlx_parallm.utils / .sample_utils / .models.base; a missing name
    is an ImportError before the server starts."""
    import sys
    import mlx_parallm_b200.compat as compat
    saved = {k: v for k, v in sys.modules.items() if k == "mlx_parallm" or k.startswith("mlx_parallm.")}
    try:
        compat.install()
        import mlx_parallm.utils as u
        for name in ("load", "generate", "stream_generate", "batch_generate", "generate_step", "batch_stream_generate_text",
                     "batch_generate_text", "apply_chat_template_cached", "encode_cached", "load_model", "load_config",
                     "get_model_path", "load_adapters", "convert", "quantize_model"):
            assert callable(getattr(u, name)), name
        from mlx_parallm.sample_utils import top_p_sampling                                    # noqa: F401
        from mlx_parallm.tokenizer_utils import TokenizerWrapper as TW, load_tokenizer        # noqa: F401
    finally:
        for k in [k for k in sys.modules if k == "mlx_parallm" or k.startswith("mlx_parallm.")]:
            del sys.modules[k]
        sys.modules.update(saved)
