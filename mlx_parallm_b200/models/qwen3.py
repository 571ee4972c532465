"""Qwen3 graph with the reference's surface (mlx_parallm/models/qwen3.py:23-226; ModelArgs restates
mlx_lm.models.qwen3.ModelArgs, which the reference imports at qwen3.py:7-13): explicit head_dim and a
per-head RMSNorm on q and k before RoPE (qwen3.py:65-70)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Union

from ._transformer import DecoderModel
from .base import BaseModelArgs


@dataclass
class ModelArgs(BaseModelArgs):
    model_type: str
    hidden_size: int
    num_hidden_layers: int
    intermediate_size: int
    num_attention_heads: int
    rms_norm_eps: float
    vocab_size: int
    num_key_value_heads: int
    max_position_embeddings: int
    rope_theta: float
    head_dim: int
    tie_word_embeddings: bool
    rope_scaling: Optional[Dict[str, Union[float, str]]] = None
    quantization: Optional[Dict[str, int]] = None

    def __post_init__(self):
        if self.rope_scaling and self.rope_scaling.get("type", self.rope_scaling.get("rope_type")) not in (None, "default", "linear"):
            raise NotImplementedError("only plain / linear RoPE is on the hot path (Hermes-4-Qwen3-14B has rope_scaling null)")


class Model(DecoderModel):
    qk_norm = True

    @staticmethod
    def _head_dim(args) -> int:
        return args.head_dim                                             # qwen3.py:33, 218-220

    def sanitize(self, weights):
        if self.args.tie_word_embeddings:                                # qwen3.py:208-211
            weights.pop("lm_head.weight", None)
            weights.pop("lm_head.scales", None)
            weights.pop("lm_head.biases", None)
        return weights
