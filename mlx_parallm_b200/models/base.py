"""KV caches and ModelArgs base with the reference's constructor / field / method surface.

Mirrors mlx_parallm/models/base.py (reference):
  BatchedKVCache  42-90   (head_dim, n_kv_heads, batch_size=1); .keys .values .offset .step
                          .offsets .reset(bs) .update_and_fetch(k, v)
  PagedKVCache    93-149  adds per-row .offsets_list
  BaseModelArgs   151-161 from_dict drops unknown keys

B200 design: one dense device buffer [B, Hkv, Tcap, D] per layer in the model's 16-bit storage type
(the reference allocates float32, base.py:111-112 -- deliberate deviation), plus a device
int32 ``lengths[B]`` the kernels read, so that a captured decode graph stays valid while rows
advance.  The model's fused kernels (b200p_rope_kv_append / b200p_attn) write and read the buffers
directly; ``update_and_fetch`` is kept for API compatibility with callers that poke the cache.
"""
from __future__ import annotations

import inspect
from dataclasses import dataclass
from typing import List, Optional

import torch


class BatchedKVCache:
    def __init__(self, head_dim, n_kv_heads, batch_size=1):
        self.n_kv_heads = n_kv_heads
        self.head_dim = head_dim
        self.batch_size = batch_size
        self.keys: Optional[torch.Tensor] = None
        self.values: Optional[torch.Tensor] = None
        self.offset = 0
        self.step = 256
        self._offsets: List[int] = [0] * batch_size
        self._lengths: Optional[torch.Tensor] = None     # device int32 [B]; may be shared across layers
        self._lengths_shared = False

    # -- reference surface ------------------------------------------------------------------
    @property
    def offsets(self):
        """Per-row offsets (base.py:87-90 / 141-143)."""
        return list(self._offsets)

    @property
    def offsets_list(self):
        return self._offsets

    @offsets_list.setter
    def offsets_list(self, v):
        self._offsets = list(v)
        self._sync_lengths()

    def reset(self, batch_size: int | None = None):
        """base.py:52-64 / 145-148: keep buffers when the batch size is unchanged."""
        if batch_size is not None and batch_size != self.batch_size:
            self.batch_size = batch_size
            self.keys = None
            self.values = None
            if not self._lengths_shared:
                self._lengths = None
        self.offset = 0
        self._offsets = [0] * self.batch_size
        if self._lengths is not None:
            if self._lengths.numel() != self.batch_size:
                self._lengths = None if not self._lengths_shared else torch.zeros(
                    (self.batch_size,), dtype=torch.int32, device=self._lengths.device)
            else:
                self._lengths.zero_()          # shared across layers: zeroing it from every layer is idempotent

    def update_and_fetch(self, keys, values):
        """Append [B, Hkv, L, D] at each row's offset and return views up to the max end
        (base.py:66-85 / 119-140).  Compatibility path only -- the model forward does not use it."""
        B, _, L, _ = keys.shape
        assert B == self.batch_size, "KV cache batch size mismatch"
        self.ensure_capacity(max(self._offsets) + L, keys.dtype, keys.device)
        for i in range(B):
            s = self._offsets[i]
            self.keys[i, :, s:s + L] = keys[i]
            self.values[i, :, s:s + L] = values[i]
        self.advance_host(L)
        self._sync_lengths()
        m = max(self._offsets)
        return self.keys[:, :, :m], self.values[:, :, :m]

    # -- B200 internals ---------------------------------------------------------------------
    @property
    def capacity(self) -> int:
        return 0 if self.keys is None else self.keys.shape[2]

    def lengths(self, device) -> torch.Tensor:
        if self._lengths is None:
            self._lengths = torch.tensor(self._offsets, dtype=torch.int32, device=device)
        return self._lengths

    def share_lengths(self, t: torch.Tensor):
        """Let every layer of one generation read (and the step advance) a single lengths tensor."""
        self._lengths = t
        self._lengths_shared = True

    def _sync_lengths(self):
        if self._lengths is not None:
            self._lengths.copy_(torch.tensor(self._offsets, dtype=torch.int32), non_blocking=False)

    def advance_host(self, L: int):
        self._offsets = [o + L for o in self._offsets]
        self.offset = max(self._offsets) if self._offsets else 0

    def ensure_capacity(self, needed: int, dtype, device) -> bool:
        """Grow in ``step`` blocks (base.py:68-80 / 104-117).  Returns True if buffers were (re)allocated
        (a captured decode graph must then be rebuilt)."""
        cap = self.capacity
        if self.keys is not None and cap >= needed and self.keys.dtype == dtype:
            return False
        n_steps = (needed + self.step - 1) // self.step
        new_cap = n_steps * self.step
        if cap > 0 and self.keys is not None and self.keys.dtype == dtype:
            # growing a live cache: at least 1.5x, so a long generation re-copies its KV (and re-captures its decode graph)
            # O(log T) times instead of every ``step`` tokens (the reference concatenates a step-sized block each time,
            # base.py:104-117).  Only ONE layer's old + new buffers are alive at a time.
            new_cap = max(new_cap, ((cap * 3 // 2 + self.step - 1) // self.step) * self.step)
        shape = (self.batch_size, self.n_kv_heads, new_cap, self.head_dim)
        nk = torch.zeros(shape, dtype=dtype, device=device)
        nv = torch.zeros(shape, dtype=dtype, device=device)
        if self.keys is not None and cap > 0:
            keep = min(cap, max(self._offsets) if self._offsets else 0)
            if keep > 0:
                nk[:, :, :keep] = self.keys[:, :, :keep]
                nv[:, :, :keep] = self.values[:, :, :keep]
        self.keys, self.values = nk, nv
        return True


class PagedKVCache(BatchedKVCache):
    """Per-sequence offsets on one dense buffer (base.py:93-149).  In this build both classes carry
    per-row device lengths, so PagedKVCache only restates the constructor for API parity."""

    def __init__(self, head_dim, n_kv_heads, batch_size=1):
        super().__init__(head_dim, n_kv_heads, batch_size)


@dataclass
class BaseModelArgs:
    @classmethod
    def from_dict(cls, params):
        return cls(**{k: v for k, v in params.items() if k in inspect.signature(cls).parameters})
