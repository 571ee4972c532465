"""Thin torch-tensor wrappers over the C-ABI (include/b200p.h).

torch is plumbing here: device memory, the current stream, CUDA-graph capture.  Every function
launches hand-written sm_100a kernels from libb200parallm.so on ``torch.cuda.current_stream()``;
none of them has a CPU or eager-PyTorch fallback -- a CPU tensor raises.
"""
from __future__ import annotations

from typing import Optional

import os

import torch

from . import _lib
from ._lib import (DQ_F32_MULADD, EPI_ADD_RESIDUAL, EPI_SILU_MUL, EPI_STORE, EPI_STORE_F32, check, lib)

_DT = {torch.float32: _lib.F32, torch.float16: _lib.F16, torch.bfloat16: _lib.BF16}


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise TypeError(f"unsupported storage dtype {dt}") from None


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dev(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.B200PError("b200p ops need CUDA tensors: there is no CPU fallback")


def _p(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


# ---- launch accounting / optional per-kernel CUDA-event timing (bench.py, profiles/) ----------------
LAUNCHES = 0            # b200p kernel-launching calls issued from Python (captured ones count once)
PROFILER = None         # set to a dict name -> list[(start_event, end_event)] to time each call


def _call(name: str, rc_fn):
    """Run one C-ABI launch; count it; bracket it with events when a profiler is attached."""
    global LAUNCHES
    LAUNCHES += 1
    if PROFILER is None:
        check(rc_fn(), name if isinstance(name, str) else name[0])
        return
    key = name if isinstance(name, str) else f"{name[0]}[N={name[1]},K={name[2]}]"
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    check(rc_fn(), key)
    e.record()
    PROFILER.setdefault(key, []).append((s, e))


# ------------------------------------------------------------------------------------------------
def dequantize(w, scales, biases, group_size=64, bits=4, mode=DQ_F32_MULADD, out=None):
    """mx.dequantize: w int32/uint32 [N, K*bits/32] -> [N, K] in scales.dtype."""
    _dev(w, scales, biases)
    N = w.shape[0]
    K = w.shape[1] * 32 // bits
    if out is None:
        out = torch.empty((N, K), dtype=scales.dtype, device=w.device)
    _call("b200p_dequantize", lambda: lib().b200p_dequantize(_p(w), _p(scales), _p(biases), _p(out), N, K, group_size, bits,
                                 dtype_code(scales.dtype), mode, _stream()))
    return out


def quantize(w, group_size=64, bits=4):
    """mx.quantize: w [N, K] (float32 / float16 / bfloat16) -> (wq int32 [N, K*bits/32], scales, biases [N, K/gs])."""
    _dev(w)
    assert w.dim() == 2 and w.is_contiguous()
    N, K = w.shape
    wq = torch.empty((N, K * bits // 32), dtype=torch.int32, device=w.device)
    s = torch.empty((N, K // group_size), dtype=w.dtype, device=w.device)
    b = torch.empty_like(s)
    _call("b200p_quantize", lambda: lib().b200p_quantize(_p(w), _p(wq), _p(s), _p(b), N, K, group_size, bits, dtype_code(w.dtype),
                                                         _stream()))
    return wq, s, b


def embed(ids, w, scales, biases, group_size=64, bits=4, mode=DQ_F32_MULADD, out=None, out_f32=True):
    """QuantizedEmbedding.__call__: ids int32 [M] -> [M, K] (float32 carrier of dtype-rounded values)."""
    _dev(ids, w, scales, biases)
    assert ids.dtype == torch.int32 and ids.is_contiguous()
    M = ids.numel()
    V = w.shape[0]
    K = w.shape[1] * 32 // bits
    if out is None:
        out = torch.empty((M, K), dtype=torch.float32 if out_f32 else scales.dtype, device=w.device)
    _call("b200p_embed", lambda: lib().b200p_embed(_p(ids), _p(w), _p(scales), _p(biases), _p(out), M, V, K, group_size, bits,
                            dtype_code(scales.dtype), mode, 1 if out_f32 else 0, _stream()))
    return out


def qmm_workspace_bytes(M, N, K) -> int:
    return int(lib().b200p_qmm_workspace_bytes(M, N, K))


_SMALLSCALE_OK = os.environ.get("B200P_NO_SMALLSCALE", "0") != "1"      # A/B switch for the subnormal-operand dequant


def qmm(x, w, scales, biases, group_size=64, bits=4, epilogue=EPI_STORE, out=None, workspace=None,
        x_perm=False, y_perm=False, small_scale=False):
    """quantized_matmul(x, w, scales, biases, transpose=True) with a fused epilogue.

    ``x_perm`` / ``y_perm``: x arrives / the 16-bit y leaves in the PERM8 column order (include/b200p.h).
    ``small_scale``: the caller has checked max|scales| < 0.0625 (B200P_W_SMALLSCALE).

    x [M, K] (row stride = x.stride(0)); out per epilogue:
      EPI_STORE -> [M, N] x.dtype; EPI_STORE_F32 -> [M, N] float32; EPI_ADD_RESIDUAL -> out (float32
      [M, N], required) += result; EPI_SILU_MUL -> [M, N/2] x.dtype."""
    _dev(x, w, scales, biases, out)
    assert x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    N = w.shape[0]
    assert w.shape[1] * 32 // bits == K, (w.shape, K, bits)
    if out is None:
        if epilogue == EPI_STORE:
            out = torch.empty((M, N), dtype=x.dtype, device=x.device)
        elif epilogue == EPI_STORE_F32:
            out = torch.empty((M, N), dtype=torch.float32, device=x.device)
        elif epilogue == EPI_SILU_MUL:
            out = torch.empty((M, N // 2), dtype=x.dtype, device=x.device)
        else:
            raise ValueError("EPI_ADD_RESIDUAL needs the residual tensor as `out`")
    epi_flags = (epilogue | (_lib.X_PERM8 if x_perm else 0) | (_lib.Y_PERM8 if y_perm else 0)
                 | (_lib.W_SMALLSCALE if (small_scale and _SMALLSCALE_OK) else 0))
    ws_ptr, ws_bytes = (0, 0) if workspace is None else (workspace.data_ptr(), workspace.numel() * workspace.element_size())
    _call(("b200p_qmm", N, K), lambda: lib().b200p_qmm(_p(x), _p(w), _p(scales), _p(biases), _p(out), M, N, K, x.stride(0), out.stride(0),
                          group_size, bits, dtype_code(x.dtype), epi_flags, ws_ptr, ws_bytes, _stream()))
    return out


def rmsnorm(x, w, eps, out=None, out_perm=False):
    """mx.fast.rms_norm on the float32 residual stream -> w.dtype (``out_perm``: PERM8 column order for a 4-bit qmm)."""
    _dev(x, w, out)
    assert x.dtype == torch.float32 and x.is_contiguous()
    M, H = x.shape
    if out is None:
        out = torch.empty((M, H), dtype=w.dtype, device=x.device)
    _call("b200p_rmsnorm", lambda: lib().b200p_rmsnorm(_p(x), _p(w), _p(out), M, H, float(eps), dtype_code(w.dtype),
                                                      _lib.OUT_PERM8 if out_perm else 0, _stream()))
    return out


def rope_kv_append(qkv, q_out, kcache, vcache, lengths, B, L, Hq, Hkv, D, rope_base, rope_scale=1.0,
                   q_norm_w=None, k_norm_w=None, norm_eps=1e-6):
    _dev(qkv, q_out, kcache, vcache, lengths)
    assert lengths.dtype == torch.int32
    Tcap = kcache.shape[2]
    _call("b200p_rope_kv_append", lambda: lib().b200p_rope_kv_append(_p(qkv), _p(q_out), _p(kcache), _p(vcache), _p(lengths), _p(q_norm_w),
                                     _p(k_norm_w), B, L, Hq, Hkv, D, Tcap, float(rope_base), float(rope_scale),
                                     float(norm_eps), dtype_code(kcache.dtype), 1 if qkv.dtype == torch.float32 and kcache.dtype != torch.float32 else 0,
                                     _stream()))


def silu_mul(gu, out, out_perm=False):
    """silu(gate) * up from the fp32 accumulation buffer gu [M, 2*I] of the fused gate|up projection; re-zeroes gu."""
    _dev(gu, out)
    assert gu.dtype == torch.float32 and gu.stride(1) == 1 and out.stride(1) == 1
    M, I = out.shape
    _call("b200p_silu_mul", lambda: lib().b200p_silu_mul(_p(gu), _p(out), M, I, gu.stride(0), out.stride(0), dtype_code(out.dtype),
                                                         _lib.OUT_PERM8 if out_perm else 0, _stream()))
    return out


def lora_add(x, At, Bt, y, col0: int, scale: float):
    """y[:, col0:col0+N] += scale * (x @ lora_a) @ lora_b with At = lora_a^T [16, K], Bt = lora_b^T [N, 16] (rank padded)."""
    _dev(x, At, Bt, y)
    M, K = x.shape
    N = Bt.shape[0]
    assert At.shape == (16, K) and Bt.shape[1] == 16 and At.is_contiguous() and Bt.is_contiguous() and x.stride(1) == 1
    _call("b200p_lora_add", lambda: lib().b200p_lora_add(_p(x), _p(At), _p(Bt), _p(y), M, N, K, x.stride(0), y.stride(0), int(col0),
                                                         float(scale), dtype_code(x.dtype), 1 if y.dtype == torch.float32 else 0,
                                                         _stream()))
    return y


def advance_lengths(lengths, delta: int):
    _dev(lengths)
    _call("b200p_advance_lengths", lambda: lib().b200p_advance_lengths(_p(lengths), lengths.numel(), int(delta), _stream()))


def kv_overflow_count(reset: bool = False) -> int:
    """Rows b200p_rope_kv_append had to drop since the last reset because their position reached the cache capacity
    (include/b200p.h).  Synchronises the current stream: call it at generation boundaries, never under graph capture."""
    import ctypes as C
    n = C.c_uint(0)
    _lib.check(lib().b200p_kv_overflow_count(C.byref(n), 1 if reset else 0, _stream()), "b200p_kv_overflow_count")
    return int(n.value)


def attn_workspace_bytes(B, L, Hq, D) -> int:
    return int(lib().b200p_attn_workspace_bytes(B, L, Hq, D))


def attn(q, kcache, vcache, lengths, out, B, L, Hq, Hkv, D, scale, workspace=None, out_perm=False):
    _dev(q, kcache, vcache, lengths, out)
    Tcap = kcache.shape[2]
    ws_ptr, ws_bytes = (0, 0) if workspace is None else (workspace.data_ptr(), workspace.numel() * workspace.element_size())
    _call("b200p_attn", lambda: lib().b200p_attn(_p(q), _p(kcache), _p(vcache), _p(lengths), _p(out), B, L, Hq, Hkv, D, Tcap, float(scale),
                           dtype_code(q.dtype), _lib.OUT_PERM8 if out_perm else 0, ws_ptr, ws_bytes, _stream()))
    return out


def argmax(logits, out=None, probs_out=None):
    """mx.argmax(logits, -1): float32 [B, V] -> int32 [B]."""
    _dev(logits, out, probs_out)
    assert logits.dtype == torch.float32 and logits.dim() == 2 and logits.stride(1) == 1
    B, V = logits.shape
    if out is None:
        out = torch.empty((B,), dtype=torch.int32, device=logits.device)
    _call("b200p_argmax", lambda: lib().b200p_argmax(_p(logits), _p(out), _p(probs_out), B, V, logits.stride(0), _stream()))
    return out


def argmax_partial(logits, vocab_offset: int, rec_out):
    """Vocab-shard reduction for the tensor-parallel greedy head: float32 [B, V_local] -> records [B, 4]."""
    _dev(logits, rec_out)
    assert logits.dtype == torch.float32 and logits.dim() == 2 and logits.stride(1) == 1
    B, V = logits.shape
    _call("b200p_argmax_partial", lambda: lib().b200p_argmax_partial(_p(logits), _p(rec_out), B, V, logits.stride(0),
                                                                     int(vocab_offset), _stream()))
    return rec_out


def argmax_merge(recs, out, probs_out=None):
    """records [W, B, 4] (all-gathered) -> tokens int32 [B] (+ softmax prob of the token)."""
    _dev(recs, out, probs_out)
    W, B = recs.shape[0], recs.shape[1]
    _call("b200p_argmax_merge", lambda: lib().b200p_argmax_merge(_p(recs), _p(out), _p(probs_out), W, B, _stream()))
    return out


def sample_top_p(logits, temperature, top_p, seed, offset, out=None, probs_out=None):
    _dev(logits, out, probs_out)
    assert logits.dtype == torch.float32 and logits.dim() == 2 and logits.stride(1) == 1
    B, V = logits.shape
    if out is None:
        out = torch.empty((B,), dtype=torch.int32, device=logits.device)
    _call("b200p_sample_top_p", lambda: lib().b200p_sample_top_p(_p(logits), _p(out), _p(probs_out), B, V, logits.stride(0), float(temperature),
                                   float(top_p), int(seed) & (2**64 - 1), int(offset) & (2**64 - 1), _stream()))
    return out
