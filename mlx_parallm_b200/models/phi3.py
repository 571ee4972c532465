"""Phi-3 graph with the reference's surface (mlx_parallm/models/phi3.py:11-214).

Same block as Llama (pre-norm residual, RoPE on q/k, GQA SDPA, silu(gate)*up); what differs is the checkpoint: q|k|v
live in ONE ``self_attn.qkv_proj`` matrix (rows [q | k | v], phi3.py:56-57, 92-96) and gate|up in ONE
``mlp.gate_up_proj`` (rows [gate | up], phi3.py:117-126) -- which is the layout the decode step wants anyway, so the
fused matrices are taken as they are and the per-projection views are cut out of them.  lm_head is always separate
(phi3.py:193).  Su-scaled RoPE (su_rope.py, the 128k-context variants) is not built: its per-frequency factors are not
an argument of b200p_rope_kv_append."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Union

from ._transformer import DecoderModel
from .base import BaseModelArgs


@dataclass
class ModelArgs(BaseModelArgs):
    """Field-for-field the reference's ModelArgs (phi3.py:11-40) plus config.json's ``quantization`` block."""
    model_type: str
    hidden_size: int
    num_hidden_layers: int
    intermediate_size: int
    num_attention_heads: int
    rms_norm_eps: float
    vocab_size: int
    num_key_value_heads: Optional[int] = None
    rope_theta: float = 10000
    rope_traditional: bool = False
    rope_scaling: Optional[Dict[str, Union[float, List[float], str]]] = None
    max_position_embeddings: int = 131072
    original_max_position_embeddings: int = 4096
    quantization: Optional[Dict[str, int]] = None

    def __post_init__(self):
        if self.num_key_value_heads is None:
            self.num_key_value_heads = self.num_attention_heads
        if self.rope_scaling:                                            # phi3.py:30-40
            if not all(k in self.rope_scaling for k in ("long_factor", "type")):
                raise ValueError("rope_scaling must contain keys {'long_factor', 'type'}")
            if self.rope_scaling["type"] == "su":
                raise NotImplementedError("Su-scaled RoPE (su_rope.py) is not built")
            if self.rope_scaling["type"] != "linear":
                self.rope_scaling = None                                 # the reference warns and drops it
        if self.rope_traditional:
            raise NotImplementedError("rope_traditional=True is not built")
        self.tie_word_embeddings = False                                 # phi3.py:193: lm_head is its own matrix


_SPLIT_LEAVES = ("weight", "scales", "biases")


class Model(DecoderModel):
    qk_norm = False

    def split_fused(self, weights: Dict) -> Dict:
        """Checkpoint keys -> the per-projection names DecoderModel.load_weights takes.  Row blocks of a fused matrix are
        views (no copy); load_weights concatenates them again into the layout they came from.  Under tensor
        parallelism tp.shard_weights must see the split names, so call this before sharding."""
        a = self.args
        hd = self._head_dim(a)
        out = {}
        for name, t in weights.items():
            mod, _, leaf = name.rpartition(".")
            if mod.endswith("self_attn.qkv_proj") and leaf in _SPLIT_LEAVES:
                nq, nkv = a.num_attention_heads * hd, a.num_key_value_heads * hd
                if t.shape[0] != nq + 2 * nkv:
                    raise ValueError(f"{name}: {t.shape[0]} rows, expected {nq + 2 * nkv}")
                base = mod[:-len("qkv_proj")]
                out[f"{base}q_proj.{leaf}"] = t[:nq]
                out[f"{base}k_proj.{leaf}"] = t[nq:nq + nkv]
                out[f"{base}v_proj.{leaf}"] = t[nq + nkv:]
            elif mod.endswith("mlp.gate_up_proj") and leaf in _SPLIT_LEAVES:
                if t.shape[0] != 2 * a.intermediate_size:
                    raise ValueError(f"{name}: {t.shape[0]} rows, expected {2 * a.intermediate_size}")
                base = mod[:-len("gate_up_proj")]
                out[f"{base}gate_proj.{leaf}"] = t[:a.intermediate_size]
                out[f"{base}up_proj.{leaf}"] = t[a.intermediate_size:]
            else:
                out[name] = t
        return out

    def sanitize(self, weights):
        return self.split_fused(super().sanitize(weights))

    def load_weights(self, weights, device="cuda", strict: bool = True):
        return super().load_weights(self.split_fused(weights), device=device, strict=strict)
