"""Nucleus sampling with the reference's signature (mlx_parallm/sample_utils.py:3-39)."""
from __future__ import annotations

import itertools

import torch

from . import ops

_counter = itertools.count()
_seed = 0


def seed(s: int) -> None:
    """Stand-in for mx.random.seed: keys the Philox stream of the sampler kernel."""
    global _seed, _counter
    _seed = int(s)
    _counter = itertools.count()


def next_offset() -> int:
    return next(_counter)


def current_seed() -> int:
    return _seed


def top_p_sampling(logits: torch.Tensor, top_p: float, temperature: float, axis: int = -1) -> torch.Tensor:
    """softmax(logits/temperature) -> keep the sorted prefix whose inclusive cumulative mass <= top_p ->
    renormalise -> draw (sample_utils.py:15-37).  One kernel (b200p_sample_top_p): radix-select of the
    nucleus threshold + inverse-CDF draw, no sort.  Returns token ids [B, 1] like the reference.
    Deviation: if even the most likely token exceeds top_p the reference divides 0/0; here the
    argmax is returned."""
    assert axis in (-1, logits.dim() - 1)
    l2 = logits.reshape(-1, logits.shape[-1]).float().contiguous()
    tok = ops.sample_top_p(l2, temperature, top_p, current_seed(), next_offset())
    return tok.to(torch.int64).reshape(*logits.shape[:-1], 1)
