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
        fx = getattr(self.model, "_tp_fused", None)
        if fx is not None and B <= fx.REC_ROWS:
            ops.argmax_partial(logits, self.v0, fx.records[:B])       # peer-visible records, gathered over NVLink
            fx.allgather_records(B, allr)
        else:
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


class _RawCuda:
    """__cuda_array_interface__ carrier so torch can view memory this library allocated (b200p_symm_alloc)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3,
                                         "strides": None}


class PeerExchange:
    """This rank's peer-mapped buffers and the fused exchange launches (csrc/allreduce.cu).

    One device allocation per rank, mapped by every peer through cudaIpc handles shipped with
    ``dist.all_gather_object`` (host plumbing, once, at attach time):

        [ partial A : max_rows x hidden f32 | partial B : same | records : 4096 x 16 B | flags | x : max_rows x hidden ]

    ``partial(i, M)`` is the tensor a row-parallel projection accumulates into (``EPI_ADD_RESIDUAL``);
    ``exchange(i, ...)`` sums the W partials over NVLink in rank order, adds the residual stream, applies the next
    RMSNorm and zeroes partial ``1 - i`` -- one kernel per rank instead of memset + ncclAllReduce + rmsnorm."""

    REC_ROWS = 4096

    def __init__(self, dist, group, rank: int, world: int, device, max_rows: int, hidden: int, mode: Optional[str] = None):
        import os
        # "rows": two-shot, row-sharded residual stream, normalised rows pushed to every rank's x (1/W of the traffic);
        # "oneshot": every rank pulls all W partials and keeps the whole residual stream
        # measured (profiles/r1_run10_*): TP=2 one-shot 9.2 us vs rows 13.8 us at M=64; TP=8 23.7 us both at M=64 but
        # rows 175 us vs one-shot 383 us at M=2048 and 2.93 vs 3.05 ms per decode step.  TP=4 (profiles/r1_run40_*): the
        # B=64 decode step is faster one-shot (Qwen3-14B 3.69 vs 4.33 ms, Llama-3-8B at 8k context 5.33 vs 5.57 ms)
        # -> rows from TP=8 up
        self.mode = mode or os.environ.get("B200P_TP_EXCHANGE", "auto")
        if self.mode == "auto":
            self.mode = "rows" if world >= 8 else "oneshot"
        assert self.mode in ("rows", "oneshot"), self.mode
        import ctypes as C
        from ._lib import check, lib
        self.rank, self.world, self.hidden, self.max_rows = rank, world, hidden, max_rows
        self.device = torch.device(device)
        L = lib()
        part = max_rows * hidden * 4
        rec = self.REC_ROWS * 16
        flag_bytes = int(L.b200p_allreduce_flag_bytes())
        self._off = (0, part, 2 * part, 2 * part + rec, 2 * part + rec + flag_bytes)
        total = 2 * part + rec + flag_bytes + part            # x region sized for float32 models too
        with torch.cuda.device(self.device):
            base = C.c_void_p()
            check(L.b200p_symm_alloc(C.byref(base), total), "b200p_symm_alloc")
            self._base = base.value
            handle = (C.c_char * 64)()
            check(L.b200p_ipc_export(self._base, handle), "b200p_ipc_export")
            handles = [None] * world
            dist.all_gather_object(handles, bytes(handle.raw), group=group)
            self._peer_base = []
            for p in range(world):
                if p == rank:
                    self._peer_base.append(self._base)
                else:
                    ptr = C.c_void_p()
                    check(L.b200p_ipc_import(handles[p], C.byref(ptr)), "b200p_ipc_import")
                    self._peer_base.append(ptr.value)
        arr = C.c_void_p * world
        self._tab = [arr(*[b + off for b in self._peer_base]) for off in self._off]      # A, B, records, flags
        self._local = [torch.as_tensor(_RawCuda(self._base + self._off[i], (max_rows, hidden), "<f4"), device=self.device)
                       for i in (0, 1)]
        self.records = torch.as_tensor(_RawCuda(self._base + self._off[2], (self.REC_ROWS, 4), "<f4"), device=self.device)
        self._x_views = {}
        self._closed = False
        dist.barrier(group=group)          # every rank has mapped every buffer before anybody launches

    def partial(self, which: int, M: int) -> torch.Tensor:
        if M > self.max_rows:
            raise ValueError(f"{M} rows exceed the peer buffer ({self.max_rows})")
        return self._local[which][:M]

    def exchange(self, which: int, h: torch.Tensor, norm_w: Optional[torch.Tensor], y: Optional[torch.Tensor], eps: float,
                 out_perm: bool = False):
        from . import _lib, ops
        M, H = h.shape
        assert H == self.hidden and h.dtype == torch.float32 and h.is_contiguous()
        dt = ops.dtype_code(norm_w.dtype) if norm_w is not None else _lib.F32
        ops._call("b200p_allreduce_residual_rmsnorm", lambda: _lib.lib().b200p_allreduce_residual_rmsnorm(
            self._tab[which], self._tab[3], self.world, self.rank, h.data_ptr(),
            0 if norm_w is None else norm_w.data_ptr(), 0 if y is None else y.data_ptr(),
            self._local[1 - which].data_ptr(), M, H, self.hidden, float(eps), dt,
            _lib.OUT_PERM8 if (out_perm and y is not None) else 0, ops._stream()))

    def x_view(self, M: int, dtype: torch.dtype) -> torch.Tensor:
        """[M, hidden] view of the peer-mapped GEMM-input buffer (what exchange_rows fills on every rank)."""
        if M > self.max_rows:
            raise ValueError(f"{M} rows exceed the peer buffer ({self.max_rows})")
        if dtype not in self._x_views:
            ts = {torch.float32: "<f4", torch.float16: "<f2", torch.bfloat16: "<i2"}[dtype]
            t = torch.as_tensor(_RawCuda(self._base + self._off[4], (self.max_rows, self.hidden), ts), device=self.device)
            self._x_views[dtype] = t.view(torch.bfloat16) if dtype == torch.bfloat16 else t
        return self._x_views[dtype][:M]

    def exchange_rows(self, which: int, h: torch.Tensor, norm_w: torch.Tensor, eps: float, out_perm: bool = False):
        """Two-shot exchange: afterwards x_view(M, norm_w.dtype) holds rms_norm(h + sum of partials) on EVERY rank and
        this rank's h holds the updated rows it owns (rows [rank*ceil(M/W), ...))."""
        from . import _lib, ops
        M, H = h.shape
        assert H == self.hidden and h.dtype == torch.float32 and h.is_contiguous()
        ops._call("b200p_allreduce_rows_rmsnorm", lambda: _lib.lib().b200p_allreduce_rows_rmsnorm(
            self._tab[which], self._tab[4], self._tab[3], self.world, self.rank, h.data_ptr(), norm_w.data_ptr(),
            self._local[1 - which].data_ptr(), M, H, self.hidden, float(eps), ops.dtype_code(norm_w.dtype),
            _lib.OUT_PERM8 if out_perm else 0, ops._stream()))

    def allgather_records(self, B: int, dst: torch.Tensor):
        """dst [W, B, 4] <- every rank's records[:B] (written by b200p_argmax_partial)."""
        from . import _lib, ops
        assert B <= self.REC_ROWS and dst.is_contiguous()
        ops._call("b200p_peer_allgather", lambda: _lib.lib().b200p_peer_allgather(
            self._tab[2], self._tab[3], self.world, self.rank, dst.data_ptr(), B * 16, ops._stream()))

    def status(self) -> int:
        import ctypes as C
        from . import _lib, ops
        out = C.c_uint32(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200p_peer_status(self._base + self._off[3], C.byref(out), ops._stream()), "b200p_peer_status")
        return int(out.value)

    def close(self):
        if self._closed:
            return
        self._closed = True
        try:                               # best effort: at interpreter shutdown the import machinery may already be gone
            from ._lib import lib
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                for p, b in enumerate(self._peer_base):
                    if p != self.rank:
                        lib().b200p_ipc_close(b)
                lib().b200p_symm_free(self._base)
        except Exception:
            pass

    def __del__(self):
        self.close()


def attach(model, dist, rank: int, world: int, group=None, allreduce: Optional[Callable] = None, fused: Optional[bool] = None):
    """Wire a model whose weights are this rank's shards (shard_weights) to the process group:
    one all-reduce(sum) of the fp32 residual stream after o_proj and after down_proj, and the
    vocab-parallel sampling head.  ``allreduce(t)`` may be supplied (e.g. the fused NVLink kernel);
    the default on GPUs is the fused peer-memory exchange (``PeerExchange``: all-reduce + residual + next RMSNorm in
    one kernel, ``fused=True``); ``fused=False`` / ``B200P_TP_FUSED=0`` keeps ``dist.all_reduce`` (NCCL ring/NVLS on
    the GPUs, gloo in CPU tests) as the baseline."""
    if world == 1:
        return model
    import os
    if fused is None:
        fused = (allreduce is None and torch.device(model.device).type == "cuda" and world in (2, 4, 8)
                 and os.environ.get("B200P_TP_FUSED", "1") != "0")
    if allreduce is None:
        def allreduce(t):
            dist.all_reduce(t, group=group)
    # communicator bring-up outside any graph capture
    probe = torch.zeros((8,), dtype=torch.float32, device=model.device)
    allreduce(probe)
    model.set_tensor_parallel(rank, world, allreduce)
    old = getattr(model, "_tp_fused", None)
    if old is not None:
        old.close()
    model._tp_fused = None
    if fused:
        from .models._transformer import MAX_TOKENS_PER_FORWARD
        model._tp_fused = PeerExchange(dist, group, rank, world, model.device, MAX_TOKENS_PER_FORWARD, model.args.hidden_size)
    model._arenas.clear()          # arenas bind x to the peer-mapped buffer in "rows" mode
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
