"""numpy restatement of the reference's model graph, KV caches and generate_step.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED vs real MLX.

Follows (paths relative to /root/reference):
  mlx_parallm/models/llama.py:49-271   Attention / MLP / TransformerBlock / LlamaModel / Model
  mlx_parallm/models/qwen3.py:23-226   same with per-head q_norm/k_norm before RoPE (65-70)
  mlx_parallm/models/base.py:42-149    BatchedKVCache / PagedKVCache
  mlx_parallm/utils.py:315-427         generate_step (greedy + sampling dispatch)

``numerics``:
  "reference"  MLX type promotion as the reference actually runs: storage dtype T up
               to the first cache fetch, PagedKVCache holds float32 (base.py:111-112), SDPA output
               and everything after it is float32.
  "mirror"     the rounding points of the CUDA path (GEMM inputs / KV in T, fp32 residual stream);
               used to separate kernel bugs from the documented B3 deviation.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

from . import mlx_ops as ops


@dataclass
class OracleConfig:
    model_type: str
    hidden_size: int
    num_hidden_layers: int
    intermediate_size: int
    num_attention_heads: int
    num_key_value_heads: int
    rms_norm_eps: float
    vocab_size: int
    head_dim: Optional[int] = None
    rope_theta: float = 10000.0
    rope_scale: float = 1.0
    tie_word_embeddings: bool = True      # llama.py:30 default
    group_size: int = 64
    bits: int = 4
    dtype: str = "float16"

    def __post_init__(self):
        if self.head_dim is None:
            self.head_dim = self.hidden_size // self.num_attention_heads   # llama.py:55

    @classmethod
    def from_hf(cls, cfg: dict, dtype: str) -> "OracleConfig":
        q = cfg.get("quantization", {})
        scale = 1.0
        rs = cfg.get("rope_scaling")
        if rs and (rs.get("type") == "linear" or rs.get("rope_type") == "linear"):
            scale = 1.0 / rs["factor"]                                      # llama.py:69-76
        mt = {"mistral": "llama"}.get(cfg["model_type"], cfg["model_type"])  # utils.py:33-36
        tie = cfg.get("tie_word_embeddings", True)
        if mt == "phi3":
            # phi3.py:43-126, 186-200: the Llama block with fused qkv_proj / gate_up_proj matrices (see split_phi3_weights)
            # and an lm_head that is never tied
            mt, tie = "llama", False
        return cls(model_type=mt, hidden_size=cfg["hidden_size"],
                   num_hidden_layers=cfg["num_hidden_layers"],
                   intermediate_size=cfg["intermediate_size"],
                   num_attention_heads=cfg["num_attention_heads"],
                   num_key_value_heads=cfg.get("num_key_value_heads") or cfg["num_attention_heads"],
                   rms_norm_eps=cfg["rms_norm_eps"], vocab_size=cfg["vocab_size"],
                   head_dim=cfg.get("head_dim") if mt == "qwen3" else None,
                   rope_theta=cfg.get("rope_theta", 10000.0), rope_scale=scale,
                   tie_word_embeddings=tie,
                   group_size=q.get("group_size", 64), bits=q.get("bits", 4), dtype=dtype)


def split_phi3_weights(weights: dict, cfg: dict) -> dict:
    """Phi-3 checkpoints keep q|k|v in one ``qkv_proj`` and gate|up in one ``gate_up_proj``; the graph splits the
    OUTPUT of those matmuls (phi3.py:92-96 ``mx.split(qkv, [Hq*D, Hq*D + Hkv*D])``, phi3.py:123-125 ``mx.split(x, 2)``),
    which equals three / two matmuls with the row blocks -- the form OracleModel evaluates."""
    D = cfg["hidden_size"] // cfg["num_attention_heads"]                    # phi3.py:53
    nq = cfg["num_attention_heads"] * D
    nkv = (cfg.get("num_key_value_heads") or cfg["num_attention_heads"]) * D
    inter = cfg["intermediate_size"]
    out = {}
    for name, t in weights.items():
        mod, _, leaf = name.rpartition(".")
        if mod.endswith("self_attn.qkv_proj"):
            base = mod[:-len("qkv_proj")]
            out[f"{base}q_proj.{leaf}"], out[f"{base}k_proj.{leaf}"], out[f"{base}v_proj.{leaf}"] = \
                t[:nq], t[nq:nq + nkv], t[nq + nkv:]
        elif mod.endswith("mlp.gate_up_proj"):
            base = mod[:-len("gate_up_proj")]
            out[f"{base}gate_proj.{leaf}"], out[f"{base}up_proj.{leaf}"] = t[:inter], t[inter:]
        else:
            out[name] = t
    return out


class OraclePagedKVCache:
    """models/base.py:93-149 -- dense [B, Hkv, Tcap, D] buffers, per-row offsets, growth in
    ``step`` blocks, float32 storage on first allocation (111-112) unless kv_dtype overrides."""

    def __init__(self, head_dim, n_kv_heads, batch_size=1, kv_dtype="float32"):
        self.head_dim, self.n_kv_heads, self.batch_size = head_dim, n_kv_heads, batch_size
        self.keys = None
        self.values = None
        self.offset = 0
        self.step = 256
        self.offsets_list = [0] * batch_size
        self.kv_dtype = kv_dtype

    @property
    def offsets(self):
        return list(self.offsets_list)

    def reset(self, batch_size=None):
        if batch_size is not None and batch_size != self.batch_size:
            self.batch_size = batch_size
            self.keys = self.values = None
        self.offset = 0
        self.offsets_list = [0] * self.batch_size

    def _ensure(self, need):
        prev = self.keys.shape[2] if self.keys is not None else 0
        if prev >= need:
            return
        n_steps = (need - prev + self.step - 1) // self.step
        shape = (self.batch_size, self.n_kv_heads, n_steps * self.step, self.head_dim)
        nk, nv = np.zeros(shape, np.float32), np.zeros(shape, np.float32)
        if self.keys is None:
            self.keys, self.values = nk, nv
        else:
            self.keys = np.concatenate([self.keys, nk], 2)
            self.values = np.concatenate([self.values, nv], 2)

    def update_and_fetch(self, k, v):
        B, _, L, _ = k.shape
        assert B == self.batch_size
        new = [o + L for o in self.offsets_list]
        mx_need = max(new)
        self._ensure(mx_need)
        for i in range(B):
            s = self.offsets_list[i]
            self.keys[i, :, s:s + L] = ops.round_to(k[i], self.kv_dtype)
            self.values[i, :, s:s + L] = ops.round_to(v[i], self.kv_dtype)
            self.offsets_list[i] = s + L
        return self.keys[:, :, :mx_need], self.values[:, :, :mx_need]


class OracleModel:
    """Holds MLX-layout weights (numpy) and evaluates the reference graph."""

    def __init__(self, cfg: OracleConfig, weights: Dict[str, np.ndarray], numerics: str = "reference",
                 tp_allreduce=None, tp_gather=None):
        """``tp_allreduce(partial) -> sum over ranks`` / ``tp_gather(local_logits) -> full logits`` turn this
        into ONE RANK of the tensor-parallel decomposition (mlx_parallm_b200/tp.py: weights are the rank's
        shards, cfg its local head counts); used by the world_size-2 gloo tests.  "mirror" numerics only."""
        self.cfg = cfg
        self.w = weights
        assert numerics in ("reference", "mirror")
        assert tp_allreduce is None or numerics == "mirror", "the TP decomposition is checked in mirror numerics"
        self.numerics = numerics
        self.tp_allreduce, self.tp_gather = tp_allreduce, tp_gather
        self._deq_cache: Dict[str, np.ndarray] = {}
        # LoRA adapters (mlx_lm LoRALinear, A.9): prefix -> (lora_a [K, r], lora_b [r, N], scale)
        self.lora: Dict[str, tuple] = {}

    # -- helpers ---------------------------------------------------------------------------
    @property
    def n_layers(self):
        return self.cfg.num_hidden_layers

    def make_cache(self, batch_size: int) -> List[OraclePagedKVCache]:
        kvd = "float32" if self.numerics == "reference" else self.cfg.dtype
        return [OraclePagedKVCache(self.cfg.head_dim, self.cfg.num_key_value_heads, batch_size, kvd)
                for _ in range(self.n_layers)]

    def _f32(self, name):
        a = self.w[name]
        return a.astype(np.float32) if a.dtype != np.float32 else a

    def _wf(self, prefix):
        """float32 W = q*s+b of one QuantizedLinear (cached; see mlx_ops.quantized_matmul)."""
        if prefix not in self._deq_cache:
            c = self.cfg
            q = ops.unpack_codes(self.w[prefix + ".weight"], c.bits).astype(np.float32)
            s = np.repeat(self._f32(prefix + ".scales"), c.group_size, -1)
            b = np.repeat(self._f32(prefix + ".biases"), c.group_size, -1)
            self._deq_cache[prefix] = q * s + b
        return self._deq_cache[prefix]

    def _linear(self, prefix, x, out_dtype):
        x = np.asarray(x, np.float32)
        y = x @ self._wf(prefix).T                                  # A.3, fp32 accumulate
        y = ops.round_to(y, out_dtype) if out_dtype else y
        if prefix in self.lora:                                     # y + (scale * (x @ A) @ B).astype(x.dtype)   (A.9)
            a, b, scale = self.lora[prefix]
            T = self.cfg.dtype
            z = ops.round_to(ops.round_to(x @ np.asarray(a, np.float32), T) @ np.asarray(b, np.float32), T)
            y = y + ops.round_to(np.float32(scale) * z, T)
            y = ops.round_to(y, out_dtype) if out_dtype else y
        return y

    # -- forward ---------------------------------------------------------------------------
    def forward(self, inputs: np.ndarray, cache: Optional[list] = None, last_only: bool = False) -> np.ndarray:
        c = self.cfg
        T = c.dtype
        ids = np.asarray(inputs)
        B, L = ids.shape
        # embed_tokens = QuantizedEmbedding -> dequantize(w[ids], ...)   (llama.py:212, A.3)
        e = "model.embed_tokens"
        h = ops.dequantize(self.w[e + ".weight"][ids.reshape(-1)], self._f32(e + ".scales")[ids.reshape(-1)],
                           self._f32(e + ".biases")[ids.reshape(-1)], c.group_size, c.bits, T).reshape(B, L, -1)
        mask = None
        if L > 1:                                                       # llama.py:214-223
            offs = cache[0].offsets if cache is not None else [0] * B
            mask = ops.additive_causal_mask_variable(L, offs, int(max(offs)) + L)
        cur = T            # dtype MLX would be carrying for activations
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        for i in range(c.num_hidden_layers):
            p = f"model.layers.{i}."
            kv = cache[i] if cache is not None else None
            x = ops.rms_norm(h, self._f32(p + "input_layernorm.weight"), c.rms_norm_eps, cur)
            if self.numerics == "mirror":
                x = ops.round_to(x, T)
            qdt = cur if self.numerics == "reference" else T
            q = self._linear(p + "self_attn.q_proj", x, qdt).reshape(B, L, H, D)
            k = self._linear(p + "self_attn.k_proj", x, qdt).reshape(B, L, Hkv, D)
            v = self._linear(p + "self_attn.v_proj", x, qdt).reshape(B, L, Hkv, D)
            if c.model_type == "qwen3":                                  # qwen3.py:65-70
                q = ops.rms_norm(q, self._f32(p + "self_attn.q_norm.weight"), c.rms_norm_eps, qdt)
                k = ops.rms_norm(k, self._f32(p + "self_attn.k_norm.weight"), c.rms_norm_eps, qdt)
            q, k, v = (t.transpose(0, 2, 1, 3) for t in (q, k, v))
            offs = kv.offsets if kv is not None else 0                   # llama.py:99-124
            q = ops.rope(q, offs, c.rope_theta, c.rope_scale, qdt)
            k = ops.rope(k, offs, c.rope_theta, c.rope_scale, qdt)
            if kv is not None:
                kf, vf = kv.update_and_fetch(k, v)
            else:
                kf, vf = k, v
            if kv is not None and self.numerics == "reference":
                cur = "float32"                                          # SURVEY B3 promotion
            odt = cur if self.numerics == "reference" else T
            o = ops.sdpa(q, kf, vf, D ** -0.5, mask, odt)
            o = o.transpose(0, 2, 1, 3).reshape(B, L, H * D)
            r = self._linear(p + "self_attn.o_proj", o, cur if self.numerics == "reference" else None)
            if self.tp_allreduce is not None:
                r = self.tp_allreduce(r)
            h = (h + r).astype(np.float32) if self.numerics == "mirror" else ops.round_to(h + r, cur)
            x = ops.rms_norm(h, self._f32(p + "post_attention_layernorm.weight"), c.rms_norm_eps,
                             cur if self.numerics == "reference" else T)
            if self.numerics == "reference":
                g = self._linear(p + "mlp.gate_proj", x, cur)
                u = self._linear(p + "mlp.up_proj", x, cur)
                a = ops.round_to(ops.round_to(ops.silu(g), cur) * u, cur)
                r = self._linear(p + "mlp.down_proj", a, cur)
                h = ops.round_to(h + r, cur)
            else:
                g = self._linear(p + "mlp.gate_proj", x, None)
                u = self._linear(p + "mlp.up_proj", x, None)
                a = ops.round_to(ops.silu(g) * u, T)
                r = self._linear(p + "mlp.down_proj", a, None)
                if self.tp_allreduce is not None:
                    r = self.tp_allreduce(r)
                h = (h + r).astype(np.float32)
        if last_only:
            h = h[:, -1:, :]
        x = ops.rms_norm(h, self._f32("model.norm.weight"), c.rms_norm_eps,
                         cur if self.numerics == "reference" else T)
        head = "model.embed_tokens" if c.tie_word_embeddings else "lm_head"   # llama.py:249-252
        logits = self._linear(head, x, cur if self.numerics == "reference" else None)
        if self.tp_gather is not None and not c.tie_word_embeddings:
            logits = self.tp_gather(logits)
        return logits.astype(np.float32)


def generate_step(prompts: np.ndarray, model: OracleModel, temp: float = 0.0, cache=None):
    """utils.py:315-427 restated for the deterministic (greedy) branch: prefill on the whole prompt,
    then one token per step; yields (tokens [B,1] int64, logits [B,V] float32 of that step).
    Sampling branches are covered distributionally by mlx_ops.top_p_distribution."""
    assert temp == 0.0, "oracle generate_step covers greedy; sampling is checked distributionally"
    y = np.asarray(prompts)
    if cache is None:
        cache = model.make_cache(y.shape[0])                            # utils.py:390-394
    while True:
        logits = model.forward(y, cache, last_only=True)[:, -1, :]      # utils.py:403-404
        y = ops.argmax_tokens(logits)                                   # utils.py:352
        yield y, logits
