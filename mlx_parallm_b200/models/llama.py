"""Llama / Mistral graph with the reference's surface (mlx_parallm/models/llama.py:15-271;
``mistral`` maps here through MODEL_REMAPPING, utils.py:33-36)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Union

from ._transformer import DecoderModel
from .base import BaseModelArgs


@dataclass
class ModelArgs(BaseModelArgs):
    """Field-for-field the reference's ModelArgs (llama.py:15-46) plus the ``quantization`` block of
    config.json that the reference reads separately (utils.py:679)."""
    model_type: str
    hidden_size: int
    num_hidden_layers: int
    intermediate_size: int
    num_attention_heads: int
    rms_norm_eps: float
    vocab_size: int
    num_key_value_heads: Optional[int] = None
    attention_bias: bool = False
    mlp_bias: bool = False
    rope_theta: float = 10000
    rope_traditional: bool = False
    rope_scaling: Optional[Dict[str, Union[float, str]]] = None
    tie_word_embeddings: bool = True
    quantization: Optional[Dict[str, int]] = None

    def __post_init__(self):
        if self.num_key_value_heads is None:
            self.num_key_value_heads = self.num_attention_heads
        if self.rope_scaling:                                            # llama.py:36-46
            if "factor" not in self.rope_scaling:
                raise ValueError("rope_scaling must contain 'factor'")
            if "type" in self.rope_scaling:
                if self.rope_scaling["type"] != "linear":
                    raise ValueError("rope_scaling 'type' currently only supports 'linear'")
            elif "rope_type" in self.rope_scaling:
                if self.rope_scaling["rope_type"] not in ["llama3", "linear"]:
                    raise ValueError(f"rope_scaling 'rope_type' {self.rope_scaling['rope_type']} not supported")
        if self.attention_bias or self.mlp_bias:
            raise NotImplementedError("projection biases are not on the hot path of any BASELINE config")
        if self.rope_traditional:
            raise NotImplementedError("rope_traditional=True is not used by Llama/Mistral checkpoints")


class Model(DecoderModel):
    qk_norm = False
