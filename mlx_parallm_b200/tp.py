"""Tensor parallelism for the decode step (new work mandated by the north star; the reference is
single-device).

One process per GPU (torchrun), ranks share a ``torch.distributed`` group (NCCL on the GPUs, gloo in
the CPU tests).  Partitioning, applied to MLX-layout tensors at load time so no re-quantisation is
needed:

    q_proj k_proj v_proj gate_proj up_proj   column-parallel: rows (output features) split; q/k/v
                                             by whole heads, so each rank owns Hq/W query heads and
                                             Hkv/W kv heads and its own KV cache shard
    o_proj down_proj                         row-parallel: the contraction axis K is split on
                                             group-size boundaries -> packed words [N, K/W*bits/32],
                                             scales/biases [N, K/W/group]; partial sums are
                                             all-reduced (one all-reduce of [B, h] fp32 after each)
    lm_head                                  vocab-parallel: rows split; greedy sampling reduces each
                                             shard to a 16-byte record per row, all-gathers the
                                             records and merges (b200p_argmax_partial / _merge);
                                             temperature sampling all-gathers the logits
    embed_tokens, norms                      replicated (tied heads therefore stay replicated)

Rank 0 carries the residual stream into each row-parallel projection's ADD_RESIDUAL epilogue, the other
ranks start from zero, so the all-reduce yields residual + full projection on every rank.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

COLUMN_PARALLEL = ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "mlp.gate_proj", "mlp.up_proj")
ROW_PARALLEL = ("self_attn.o_proj", "mlp.down_proj")


def _head_dim(cfg: dict) -> int:
    return cfg.get("head_dim") or cfg["hidden_size"] // cfg["num_attention_heads"]


def check_divisible(cfg: dict, world: int) -> None:
    """Raise ValueError unless every split falls on a head / quantisation-group boundary."""
    Hq = cfg["num_attention_heads"]
    Hkv = cfg.get("num_key_value_heads") or Hq
    gs = cfg["quantization"]["group_size"]
    inter = cfg["intermediate_size"]
    if world < 1:
        raise ValueError("world must be >= 1")
    if Hq % world or Hkv % world:
        raise ValueError(f"heads ({Hq} q / {Hkv} kv) are not divisible by TP={world}")
    if (Hq // world * _head_dim(cfg)) % gs:
        raise ValueError("o_proj K shard does not fall on a quantisation-group boundary")
    if inter % world or (inter // world) % gs:
        raise ValueError(f"intermediate_size {inter} / TP={world} does not fall on a group-{gs} boundary")
    if not cfg.get("tie_word_embeddings", True) and (cfg["vocab_size"] % world or (cfg["vocab_size"] // world) % 2):
        raise ValueError(f"vocab_size {cfg['vocab_size']} is not divisible into even shards by TP={world}")


def vocab_range(cfg: dict, rank: int, world: int):
    """[start, end) of the lm_head rows held by ``rank`` (whole vocab for tied heads)."""
    V = cfg["vocab_size"]
    if cfg.get("tie_word_embeddings", True) or world == 1:
        return 0, V
    per = V // world
    return rank * per, (rank + 1) * per


def shard_weights(weights: Dict[str, torch.Tensor], cfg: dict, rank: int, world: int) -> Dict[str, torch.Tensor]:
    """Return this rank's shard of an MLX-layout state dict (views/copies of the input tensors)."""
    if world == 1:
        return dict(weights)
    check_divisible(cfg, world)
    bits = cfg["quantization"]["bits"]
    gs = cfg["quantization"]["group_size"]
    out: Dict[str, torch.Tensor] = {}
    for name, t in weights.items():
        mod, _, leaf = name.rpartition(".")
        if mod.endswith(COLUMN_PARALLEL) or (mod == "lm_head"):
            n = t.shape[0]
            assert n % world == 0, (name, n, world)
            per = n // world
            out[name] = t[rank * per:(rank + 1) * per].contiguous()
        elif mod.endswith(ROW_PARALLEL):
            cols = t.shape[1]
            assert cols % world == 0, (name, cols, world)
            per = cols // world
            # weight: K*bits/32 words per row; scales/biases: K/gs per row -- both split evenly iff K/W % gs == 0
            if leaf == "weight":
                assert (per * 32 // bits) % gs == 0, (name, per)
            out[name] = t[:, rank * per:(rank + 1) * per].contiguous()
        else:
            out[name] = t
    return out


def local_config(cfg: dict, world: int) -> dict:
    """The per-rank view of an HF config (head counts, intermediate size, vocab shard) -- what an
    unsharded model of the shard's shape would be configured with."""
    c = dict(cfg)
    c["head_dim"] = _head_dim(cfg)
    c["num_attention_heads"] = cfg["num_attention_heads"] // world
    c["num_key_value_heads"] = (cfg.get("num_key_value_heads") or cfg["num_attention_heads"]) // world
    c["intermediate_size"] = cfg["intermediate_size"] // world
    return c


# --------------------------------------------------------------------------------------------- runtime
class TensorParallelHead:
    """Vocab-parallel sampling head bound to a model: replaces DecoderModel.greedy_tokens / sample_tokens."""

    def __init__(self, model, dist, group, rank: int, world: int, vocab_start: int, sharded_vocab: bool):
        self.model, self.dist, self.group = model, dist, group
        self.rank, self.world, self.v0, self.sharded = rank, world, vocab_start, sharded_vocab
        self._rec: Dict[int, tuple] = {}
        self._gath: Dict[tuple, torch.Tensor] = {}

    def greedy(self, logits, tok_out, probs_out=None):
        from . import ops
        if not self.sharded:
            return ops.argmax(logits, out=tok_out, probs_out=probs_out)
        B = logits.shape[0]
        if B not in self._rec:
            dev = logits.device
            self._rec[B] = (torch.empty((B, 4), dtype=torch.float32, device=dev),
                            torch.empty((self.world, B, 4), dtype=torch.float32, device=dev))
        mine, allr = self._rec[B]
        ops.argmax_partial(logits, self.v0, mine)
        self.dist.all_gather_into_tensor(allr.view(-1), mine.view(-1), group=self.group)
        ops.argmax_merge(allr, tok_out, probs_out)

    def sample(self, logits, temp, top_p, seed, offset, tok_out, probs_out=None):
        from . import ops
        if not self.sharded:
            return ops.sample_top_p(logits, temp, top_p, seed, offset, out=tok_out, probs_out=probs_out)
        B, Vl = logits.shape
        key = (B, Vl)
        if key not in self._gath:
            self._gath[key] = torch.empty((self.world, B, Vl), dtype=torch.float32, device=logits.device)
        g = self._gath[key]
        self.dist.all_gather_into_tensor(g.view(-1), logits.contiguous().view(-1), group=self.group)
        full = g.permute(1, 0, 2).reshape(B, self.world * Vl).contiguous()
        # every rank draws with the same Philox (seed, offset): identical tokens without a broadcast
        ops.sample_top_p(full, temp, top_p, seed, offset, out=tok_out, probs_out=probs_out)


def attach(model, dist, rank: int, world: int, group=None, allreduce: Optional[Callable] = None):
    """Wire a model whose weights are this rank's shards (shard_weights) to the process group:
    one all-reduce(sum) of the fp32 residual stream after o_proj and after down_proj, and the
    vocab-parallel sampling head.  ``allreduce(t)`` may be supplied (e.g. the fused NVLink kernel);
    the default is ``dist.all_reduce`` (NCCL ring/NVLS on the GPUs, gloo in CPU tests)."""
    if world == 1:
        return model
    if allreduce is None:
        def allreduce(t):
            dist.all_reduce(t, group=group)
    # communicator bring-up outside any graph capture
    probe = torch.zeros((8,), dtype=torch.float32, device=model.device)
    allreduce(probe)
    model.set_tensor_parallel(rank, world, allreduce)
    sharded_vocab = model.lm_head is not None
    v0 = rank * model.lm_head.out_features if sharded_vocab else 0
    head = TensorParallelHead(model, dist, group, rank, world, v0, sharded_vocab)
    if sharded_vocab:
        g = torch.empty((world * 8,), dtype=torch.float32, device=model.device)
        dist.all_gather_into_tensor(g, probe, group=group)
    model.greedy_tokens = head.greedy
    model.sample_tokens = head.sample
    model._tp_head = head
    return model
