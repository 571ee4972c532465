"""Deterministic synthetic checkpoints in the reference's on-disk format.

There is no network for real checkpoints and MLX's threefry PRNG is not reproducible outside MLX
(scripts/build_tiny_model.py:123 in the reference), so every BASELINE config is instantiated from
numpy / torch generators with fixed seeds, in MLX layout: ``*.weight`` uint32 packed codes,
``*.scales`` / ``*.biases`` [N, K/64], norm weights, ``config.json`` with a ``quantization`` block
(utils.py:679), ``model.safetensors`` with metadata {"format": "mlx"} (utils.py:870).
"""
from __future__ import annotations

import json
import math
import os
from typing import Dict, Optional

import numpy as np
import torch

# HF-style configs of the five BASELINE.json configs (+ tiny variants for tests).  Dims: SURVEY 8.
CONFIGS: Dict[str, dict] = {
    # scripts/build_tiny_model.py:70-101 (hidden 64, 8 layers, 4 heads, inter 128, tied, 4-bit g64)
    "tiny": dict(model_type="llama", hidden_size=64, num_hidden_layers=8, intermediate_size=128,
                 num_attention_heads=4, num_key_value_heads=4, rms_norm_eps=1e-6, vocab_size=151936,
                 rope_theta=10000.0, tie_word_embeddings=True, quantization=dict(group_size=64, bits=4)),
    # same shape family with a small vocab so CPU tests stay fast
    "tiny-llama": dict(model_type="llama", hidden_size=128, num_hidden_layers=3, intermediate_size=256,
                       num_attention_heads=4, num_key_value_heads=2, rms_norm_eps=1e-5, vocab_size=512,
                       rope_theta=10000.0, tie_word_embeddings=False, quantization=dict(group_size=64, bits=4)),
    "tiny-qwen3": dict(model_type="qwen3", hidden_size=128, num_hidden_layers=3, intermediate_size=256,
                       num_attention_heads=4, num_key_value_heads=2, head_dim=64, rms_norm_eps=1e-6,
                       vocab_size=512, rope_theta=1000000.0, max_position_embeddings=4096,
                       tie_word_embeddings=False, quantization=dict(group_size=64, bits=4)),
    # D=128 / GQA-4 model small enough for the oracle, wide enough for the tensor paths
    "small-llama": dict(model_type="llama", hidden_size=512, num_hidden_layers=2, intermediate_size=1024,
                        num_attention_heads=4, num_key_value_heads=1, rms_norm_eps=1e-5, vocab_size=1024,
                        rope_theta=500000.0, tie_word_embeddings=False, quantization=dict(group_size=64, bits=4)),
    # Phi-3 graph (fused qkv_proj / gate_up_proj checkpoint keys, models/phi3.py); D=128, GQA-2
    "small-phi3": dict(model_type="phi3", hidden_size=512, num_hidden_layers=2, intermediate_size=1024,
                       num_attention_heads=4, num_key_value_heads=2, rms_norm_eps=1e-5, vocab_size=1024,
                       rope_theta=10000.0, tie_word_embeddings=False, quantization=dict(group_size=64, bits=4)),
    "mistral-7b": dict(model_type="mistral", hidden_size=4096, num_hidden_layers=32, intermediate_size=14336,
                       num_attention_heads=32, num_key_value_heads=8, rms_norm_eps=1e-5, vocab_size=32000,
                       rope_theta=10000.0, tie_word_embeddings=False, sliding_window=4096,
                       quantization=dict(group_size=64, bits=4)),
    "llama-3-8b": dict(model_type="llama", hidden_size=4096, num_hidden_layers=32, intermediate_size=14336,
                       num_attention_heads=32, num_key_value_heads=8, rms_norm_eps=1e-5, vocab_size=128256,
                       rope_theta=500000.0, tie_word_embeddings=False, quantization=dict(group_size=64, bits=4)),
    "qwen3-14b": dict(model_type="qwen3", hidden_size=5120, num_hidden_layers=40, intermediate_size=17408,
                      num_attention_heads=40, num_key_value_heads=8, head_dim=128, rms_norm_eps=1e-6,
                      vocab_size=151936, rope_theta=1000000.0, max_position_embeddings=40960,
                      tie_word_embeddings=False, quantization=dict(group_size=64, bits=4)),
    "llama-3-70b": dict(model_type="llama", hidden_size=8192, num_hidden_layers=80, intermediate_size=28672,
                        num_attention_heads=64, num_key_value_heads=8, rms_norm_eps=1e-5, vocab_size=128256,
                        rope_theta=500000.0, tie_word_embeddings=False, quantization=dict(group_size=64, bits=8)),
}


def get_config(name: str, **overrides) -> dict:
    cfg = json.loads(json.dumps(CONFIGS[name]))
    cfg.update(overrides)
    return cfg


def head_dim_of(cfg: dict) -> int:
    return cfg.get("head_dim") or cfg["hidden_size"] // cfg["num_attention_heads"]


def linear_shapes(cfg: dict) -> Dict[str, tuple]:
    """(N, K) of every QuantizedLinear / QuantizedEmbedding by parameter prefix."""
    h, inter, D = cfg["hidden_size"], cfg["intermediate_size"], head_dim_of(cfg)
    Hq, Hkv = cfg["num_attention_heads"], cfg.get("num_key_value_heads") or cfg["num_attention_heads"]
    out = {"model.embed_tokens": (cfg["vocab_size"], h)}
    if not cfg.get("tie_word_embeddings", True):
        out["lm_head"] = (cfg["vocab_size"], h)
    for i in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{i}."
        out[p + "self_attn.q_proj"] = (Hq * D, h)
        out[p + "self_attn.k_proj"] = (Hkv * D, h)
        out[p + "self_attn.v_proj"] = (Hkv * D, h)
        out[p + "self_attn.o_proj"] = (h, Hq * D)
        out[p + "mlp.gate_proj"] = (inter, h)
        out[p + "mlp.up_proj"] = (inter, h)
        out[p + "mlp.down_proj"] = (h, inter)
    return out


def norm_names(cfg: dict):
    names = ["model.norm.weight"]
    for i in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{i}."
        names += [p + "input_layernorm.weight", p + "post_attention_layernorm.weight"]
        if cfg["model_type"] == "qwen3":
            names += [p + "self_attn.q_norm.weight", p + "self_attn.k_norm.weight"]
    return names


def weight_bytes_per_step(cfg: dict, scale_bytes: int = 2) -> int:
    """Algorithmic weight bytes of one decode step (SURVEY 8d / BASELINE.md section 3): every linear
    except the embedding table (B gathered rows only) + norm weights."""
    q = cfg["quantization"]
    total = 0
    for name, (n, k) in linear_shapes(cfg).items():
        if name == "model.embed_tokens" and not cfg.get("tie_word_embeddings", True):
            continue
        total += n * k * q["bits"] // 8 + 2 * n * (k // q["group_size"]) * scale_bytes
    total += (2 * cfg["num_hidden_layers"] + 1) * cfg["hidden_size"] * scale_bytes
    return total


def kv_bytes_per_token(cfg: dict, kv_bytes: int = 2) -> int:
    Hkv = cfg.get("num_key_value_heads") or cfg["num_attention_heads"]
    return 2 * Hkv * head_dim_of(cfg) * cfg["num_hidden_layers"] * kv_bytes


_TORCH_DT = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def _stable_seed(seed: int, name: str) -> int:
    h = 1469598103934665603
    for ch in name.encode():
        h = ((h ^ ch) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return (h ^ (seed * 0x9E3779B97F4A7C15)) & 0x7FFFFFFFFFFFFFFF


def random_weights(cfg: dict, seed: int = 1234, dtype: str = "float16", device="cpu",
                   only: Optional[callable] = None) -> Dict[str, torch.Tensor]:
    """Config-2 style generator: uniform random codes, scales = +-U(0.5,1.5)*(2/(2^bits-1))/sqrt(K)
    (random sign exercises MLX's signed scales), biases = -scale*(2^bits-1)/2 + N(0, 0.01/sqrt(K)),
    norm weights 1 + N(0, 0.02).  One torch.Generator per tensor, seeded from (seed, name), so any subset or
    any TP shard of the same model can be regenerated independently on any device type... but NOT across
    device types (CPU and CUDA generators differ): parity tests generate on CPU and copy."""
    q = cfg["quantization"]
    bits, gs = q["bits"], q["group_size"]
    nb = float((1 << bits) - 1)
    tdt = _TORCH_DT[dtype]
    dev = torch.device(device)
    out: Dict[str, torch.Tensor] = {}
    for name, (n, k) in linear_shapes(cfg).items():
        if only is not None and not only(name):
            continue
        g = torch.Generator(device=dev)
        g.manual_seed(_stable_seed(seed, name))
        words = k * bits // 32
        w = torch.randint(-(2 ** 31), 2 ** 31, (n, words), generator=g, device=dev, dtype=torch.int64).to(torch.int32)
        mag = (torch.rand((n, k // gs), generator=g, device=dev) + 0.5) * (2.0 / nb) / math.sqrt(k)
        sign = torch.where(torch.rand((n, k // gs), generator=g, device=dev) < 0.5, -1.0, 1.0)
        s = (mag * sign).to(tdt)
        b = (-s.float() * (nb / 2) + torch.randn((n, k // gs), generator=g, device=dev) * (0.01 / math.sqrt(k))).to(tdt)
        if name == "model.embed_tokens":                   # embeddings ~ N(0, 1/sqrt(h))-ish spread
            s = (s.float() * math.sqrt(k) * 0.5).to(tdt)
            b = (-s.float() * (nb / 2)).to(tdt)
        out[name + ".weight"], out[name + ".scales"], out[name + ".biases"] = w, s, b
    for name in norm_names(cfg):
        if only is not None and not only(name):
            continue
        g = torch.Generator(device=dev)
        g.manual_seed(_stable_seed(seed, name))
        dim = head_dim_of(cfg) if ("q_norm" in name or "k_norm" in name) else cfg["hidden_size"]
        out[name] = (1.0 + 0.02 * torch.randn((dim,), generator=g, device=dev)).to(tdt)
    return out


def to_phi3_checkpoint(weights: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Per-projection tensors -> the key layout of a Phi-3 checkpoint: rows [q | k | v] as ``self_attn.qkv_proj`` and
    [gate | up] as ``mlp.gate_up_proj`` (reference phi3.py:56-57, 92-96, 117-126)."""
    out = {}
    for name, t in weights.items():
        mod, _, leaf = name.rpartition(".")
        if mod.endswith(("self_attn.k_proj", "self_attn.v_proj", "mlp.up_proj")):
            continue
        if mod.endswith("self_attn.q_proj"):
            base = mod[:-len("q_proj")]
            out[f"{base}qkv_proj.{leaf}"] = torch.cat([weights[f"{base}{p}_proj.{leaf}"] for p in "qkv"], 0)
        elif mod.endswith("mlp.gate_proj"):
            base = mod[:-len("gate_proj")]
            out[f"{base}gate_up_proj.{leaf}"] = torch.cat([t, weights[f"{base}up_proj.{leaf}"]], 0)
        else:
            out[name] = t
    return out


def to_numpy_weights(weights: Dict[str, torch.Tensor]) -> Dict[str, np.ndarray]:
    """torch -> numpy for the oracle: packed int32 -> uint32, bf16/f16 -> float32 (exact)."""
    out = {}
    for k, v in weights.items():
        v = v.detach().cpu()
        if v.dtype == torch.int32:
            out[k] = v.numpy().view(np.uint32)
        else:
            out[k] = v.float().numpy()
    return out


def save_checkpoint(path: str, cfg: dict, weights: Dict[str, torch.Tensor]):
    """Write config.json + model.safetensors the way the reference's save_weights does (utils.py:836-885)."""
    from safetensors.torch import save_file
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(cfg, f, indent=2)
    tensors = {}
    for k, v in weights.items():
        v = v.detach().cpu().contiguous()
        tensors[k] = v.view(torch.uint32) if v.dtype == torch.int32 else v
    save_file(tensors, os.path.join(path, "model.safetensors"), metadata={"format": "mlx"})
