"""numpy restatement of the MLX primitives on the decode hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED vs real MLX.

Every function cites the reference call site it stands in for (paths relative to
/root/reference) and the SURVEY.md appendix item that states the MLX semantics.

dtype convention: arrays are carried as numpy float32 (or float16) and a ``dtype``
string in {"float32", "float16", "bfloat16"} says which storage type MLX would
have used; ``round_to`` applies that type's round-to-nearest-even.
"""
from __future__ import annotations

import numpy as np

DTYPES = ("float32", "float16", "bfloat16")

# dequantize rounding modes (SURVEY Appendix A.2 -- unverifiable without MLX, so all
# candidates are implemented; the CUDA kernel must match whichever is selected, bit for bit)
MODE_F32_MULADD = 0   # p = fl32(q*s); r = fl32(p + b); out = round_T(r)      (MLX CPU-backend style)
MODE_FMA = 1          # out = round_T(fma32(q, s, b))
MODE_T_MULADD = 2     # p = round_T(q*s); out = round_T(p + b)                 (storage-type arithmetic)


# ----------------------------------------------------------------------------- rounding helpers
def bf16_round(x: np.ndarray) -> np.ndarray:
    """float32 -> nearest-even bfloat16, returned as float32 values."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    nan = np.isnan(x)
    lsb = (u >> 16) & 1
    r = ((u + 0x7FFF + lsb) >> 16) << 16
    out = (r & 0xFFFFFFFF).astype(np.uint32).view(np.float32)
    out = np.where(nan, np.float32(np.nan), out)
    return out.reshape(x.shape)


def bf16_bits(x: np.ndarray) -> np.ndarray:
    """bf16-representable float32 values -> uint16 bit patterns."""
    return (np.ascontiguousarray(x, dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)


def bf16_from_bits(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << 16).view(np.float32)


def e4m3_round(x: np.ndarray) -> np.ndarray:
    """float32 -> nearest OCP FP8 E4M3 value (round-half-even, 3 mantissa bits, normals 2^-6 .. 448, subnormals in steps
    of 2^-9, saturating), returned as float32.  Not a reference format: the reference stores K/V in float32 (base.py:111-112);
    this exists to QUANTIFY what an fp8 KV cache would cost against the 1e-2 logits budget."""
    x = np.asarray(x, dtype=np.float32)
    a = np.abs(x).astype(np.float64)
    with np.errstate(divide="ignore"):
        e = np.floor(np.log2(np.where(a > 0, a, 1.0)))
    e = np.clip(e, -6, 8)
    step = np.exp2(e - 3)                                 # spacing of the 3-bit mantissa at this exponent
    q = np.rint(a / step) * step                          # rint = half-to-even
    q = np.minimum(q, 448.0)
    return (np.sign(x) * q).astype(np.float32)


def round_to(x: np.ndarray, dtype: str) -> np.ndarray:
    """Round float32 values to ``dtype`` storage precision; result stays float32."""
    x = np.asarray(x, dtype=np.float32)
    if dtype == "float32":
        return x
    if dtype == "float8_e4m3":
        return e4m3_round(x)
    if dtype == "float16":
        with np.errstate(over="ignore"):
            return x.astype(np.float16).astype(np.float32)
    if dtype == "bfloat16":
        return bf16_round(x)
    raise ValueError(dtype)


def storage_bits(x: np.ndarray, dtype: str) -> np.ndarray:
    """Bit pattern of already-rounded values in their storage dtype (for bit-exact compares)."""
    x = np.asarray(x, dtype=np.float32)
    if dtype == "float32":
        return x.view(np.uint32)
    if dtype == "float16":
        return x.astype(np.float16).view(np.uint16)
    return bf16_bits(x)


# ----------------------------------------------------------------------------- quantize / pack
def pack_codes(q: np.ndarray, bits: int) -> np.ndarray:
    """[N, K] integer codes -> [N, K*bits/32] uint32 (A.1: element j of each run of 32/bits
    consecutive input-axis elements occupies bits [j*bits, (j+1)*bits), little end first)."""
    q = np.asarray(q).astype(np.uint32)
    n, k = q.shape
    per = 32 // bits
    assert k % per == 0
    q = q.reshape(n, k // per, per)
    shifts = (np.arange(per, dtype=np.uint32) * np.uint32(bits))
    return np.bitwise_or.reduce(q << shifts, axis=-1).astype(np.uint32)


def unpack_codes(w: np.ndarray, bits: int) -> np.ndarray:
    """inverse of pack_codes -> [N, K] uint32 codes."""
    w = np.asarray(w, dtype=np.uint32)
    per = 32 // bits
    shifts = (np.arange(per, dtype=np.uint32) * np.uint32(bits))
    q = (w[..., None] >> shifts) & np.uint32((1 << bits) - 1)
    return q.reshape(*w.shape[:-1], w.shape[-1] * per)


def quantize(w: np.ndarray, group_size: int = 64, bits: int = 4, dtype: str = "float32"):
    """mx.quantize (A.1; used by nn.quantize at utils.py:686-690 and scripts/build_tiny_model.py:151).

    Returns (w_q uint32 [N, K*bits/32], scales [N, K/gs], biases [N, K/gs]) with scales/biases
    rounded to ``dtype`` (float32 carrier)."""
    w = round_to(w, dtype)
    n, k = w.shape
    assert k % group_size == 0
    g = w.reshape(n, k // group_size, group_size).astype(np.float32)
    n_bins = np.float32((1 << bits) - 1)
    eps = np.float32(1e-7)
    w_max = g.max(-1)
    w_min = g.min(-1)
    mask = np.abs(w_min) > np.abs(w_max)
    scales = np.maximum(round_to((w_max - w_min) / n_bins, dtype), eps)
    scales = np.where(mask, scales, -scales).astype(np.float32)
    edge = np.where(mask, w_min, w_max).astype(np.float32)
    q0 = np.rint(round_to(edge / scales, dtype))
    with np.errstate(divide="ignore", invalid="ignore"):
        scales = np.where(q0 != 0, round_to(edge / np.where(q0 != 0, q0, 1), dtype), scales)
    biases = np.where(q0 == 0, np.float32(0), edge).astype(np.float32)
    scales = round_to(scales, dtype)
    biases = round_to(biases, dtype)
    q = np.rint(round_to((g - biases[..., None]) / scales[..., None], dtype))
    q = np.clip(q, 0, n_bins).astype(np.uint32).reshape(n, k)
    return pack_codes(q, bits), scales, biases


def dequantize(w_q, scales, biases, group_size: int = 64, bits: int = 4,
               dtype: str = "float32", mode: int = MODE_F32_MULADD) -> np.ndarray:
    """mx.dequantize (A.2; reached through nn.QuantizedEmbedding.__call__, llama.py:212 / qwen3.py:166).

    w[n,k] = q[n,k] * scales[n,k//gs] + biases[n,k//gs]; result rounded to ``dtype``."""
    q = unpack_codes(w_q, bits).astype(np.float32)
    n, k = q.shape
    s = np.repeat(np.asarray(scales, np.float32), group_size, axis=-1)
    b = np.repeat(np.asarray(biases, np.float32), group_size, axis=-1)
    if mode == MODE_F32_MULADD:
        p = (q * s).astype(np.float32)
        r = (p + b).astype(np.float32)
    elif mode == MODE_FMA:
        # exact in float64 (q<=255, 24-bit significands -> product exact, sum exact enough: the
        # float64 sum of an exact <=32-bit product and a 24-bit addend rounds once to f32 below)
        r = (q.astype(np.float64) * s.astype(np.float64) + b.astype(np.float64)).astype(np.float32)
    elif mode == MODE_T_MULADD:
        p = round_to(q * s, dtype)
        r = (p + b).astype(np.float32)
    else:
        raise ValueError(mode)
    return round_to(r, dtype)


def quantized_matmul(x, w_q, scales, biases, group_size: int = 64, bits: int = 4,
                     dtype: str = "float32") -> np.ndarray:
    """mx.quantized_matmul(x, w, scales, biases, transpose=True) (A.3; every nn.QuantizedLinear
    on the path: llama.py:93,143,165,250-252).  y = x @ dequant(w).T, fp32 accumulation in the
    grouped form  sum_g scale_g*(sum_k x_k q_k) + bias_g*(sum_k x_k); output rounded to ``dtype``."""
    x = np.asarray(x, np.float32)
    lead = x.shape[:-1]
    k = x.shape[-1]
    x2 = x.reshape(-1, k)
    q = unpack_codes(w_q, bits).astype(np.float32)            # [N, K]
    n = q.shape[0]
    ng = k // group_size
    s = np.asarray(scales, np.float32)                         # [N, G]
    b = np.asarray(biases, np.float32)
    # grouped form, but evaluated with one BLAS call: W = q*s + b exactly in fp32-ish
    # (sum_g s_g * (x_g . q_g)) == x . (q * s_rep); the difference to the per-group order is
    # fp32 reassociation only, far inside the 1e-2 tolerance the north star states.
    wf = q * np.repeat(s, group_size, -1) + np.repeat(b, group_size, -1)
    y = x2 @ wf.T
    del ng, n
    return round_to(y.reshape(*lead, -1), dtype)


def quantized_matmul_grouped(x, w_q, scales, biases, group_size=64, bits=4) -> np.ndarray:
    """Literal grouped evaluation order of A.3 in float64 (slow; small shapes) -- used by tests to
    bound how far any fp32 evaluation order may sit from the exact value."""
    x = np.asarray(x, np.float64)
    k = x.shape[-1]
    x2 = x.reshape(-1, k)
    q = unpack_codes(w_q, bits).astype(np.float64)
    g = k // group_size
    xg = x2.reshape(x2.shape[0], g, group_size)
    qg = q.reshape(q.shape[0], g, group_size)
    dots = np.einsum("mgk,ngk->mng", xg, qg)
    sums = xg.sum(-1)
    y = (dots * np.asarray(scales, np.float64)[None]).sum(-1) + (sums[:, None, :] * np.asarray(biases, np.float64)[None]).sum(-1)
    return y.reshape(*x.shape[:-1], -1)


# ----------------------------------------------------------------------------- norm / rope / attention
def rms_norm(x, w, eps: float, dtype: str = "float32") -> np.ndarray:
    """mx.fast.rms_norm (A.4; nn.RMSNorm at llama.py:187,189,231, qwen3.py:42-43,65-70).
    fp32: x * rsqrt(mean(x^2) + eps) * w, rounded once to ``dtype``."""
    x = np.asarray(x, np.float32)
    ms = np.mean(x.astype(np.float32) ** 2, axis=-1, keepdims=True, dtype=np.float32)
    y = x * (np.float32(1.0) / np.sqrt(ms + np.float32(eps))) * np.asarray(w, np.float32)
    return round_to(y, dtype)


def rope(x, offset, base: float, scale: float = 1.0, dtype: str = "float32") -> np.ndarray:
    """mx.fast.rope(x[B,H,L,D], dims=D, traditional=False, base, scale, offset) (A.5; nn.RoPE at
    llama.py:107-124).  ``offset`` may be an int or a per-row array [B] (the reference loops rows,
    llama.py:107-117).  Non-traditional pairing (i, i+D/2)."""
    x = np.asarray(x, np.float32)
    B, H, L, D = x.shape
    half = D // 2
    off = np.broadcast_to(np.asarray(offset, np.float32).reshape(-1), (B,)) if np.ndim(offset) else np.full((B,), offset, np.float32)
    pos = (off[:, None] + np.arange(L, dtype=np.float32)[None]) * np.float32(scale)      # [B, L]
    inv = np.exp(-np.arange(half, dtype=np.float32) * (np.log(np.float32(base)) / np.float32(half))).astype(np.float32)
    th = pos[:, None, :, None] * inv[None, None, None, :]                                 # [B,1,L,half]
    c, s = np.cos(th).astype(np.float32), np.sin(th).astype(np.float32)
    x1, x2 = x[..., :half], x[..., half:]
    out = np.concatenate([x1 * c - x2 * s, x1 * s + x2 * c], axis=-1)
    return round_to(out, dtype)


def sdpa(q, k, v, scale: float, mask=None, dtype: str = "float32") -> np.ndarray:
    """mx.fast.scaled_dot_product_attention (A.6; llama.py:139-141, qwen3.py:111-113).
    q [B,Hq,Lq,D], k/v [B,Hkv,T,D]; GQA: q head h uses kv head h // (Hq/Hkv); fp32 softmax."""
    q = np.asarray(q, np.float32)
    k = np.asarray(k, np.float32)
    v = np.asarray(v, np.float32)
    B, Hq, Lq, D = q.shape
    Hkv = k.shape[1]
    G = Hq // Hkv
    T = k.shape[2]
    qg = (q * np.float32(scale)).reshape(B, Hkv, G * Lq, D)
    s = np.matmul(qg, k.transpose(0, 1, 3, 2)).reshape(B, Hkv, G, Lq, T)
    if mask is not None:
        m = np.asarray(mask, np.float32)
        m = m.reshape(m.shape[0], 1, 1, *m.shape[-2:]) if m.ndim == 3 else m.reshape(1, 1, 1, *m.shape[-2:])
        s = s + m
    s = s - s.max(-1, keepdims=True)
    p = np.exp(s)
    p = p / p.sum(-1, keepdims=True)
    o = np.matmul(p.reshape(B, Hkv, G * Lq, T), v).reshape(B, Hq, Lq, D)
    return round_to(o, dtype)


def silu(x):
    """nn.silu (A.7; llama.py:165)."""
    x = np.asarray(x, np.float32)
    return x / (np.float32(1.0) + np.exp(-x))


def additive_causal_mask_variable(n: int, offsets, total_length: int) -> np.ndarray:
    """create_additive_causal_mask_variable (models/base.py:17-40): [B, n, total] with -1e9 above
    the causal boundary of each row (offset_b + i < j)."""
    offs = np.asarray(offsets).reshape(-1)
    r = np.arange(total_length)[None, None, :]
    l = (offs[:, None] + np.arange(n)[None, :])[:, :, None]
    return ((l < r) * np.float32(-1e9)).astype(np.float32)


# ----------------------------------------------------------------------------- sampling
def argmax_tokens(logits) -> np.ndarray:
    """mx.argmax(logits, -1, keepdims=True) (utils.py:352); lowest index among ties (A.7)."""
    return np.argmax(np.asarray(logits, np.float32), axis=-1)[:, None].astype(np.int64)


def softmax(x, axis=-1):
    x = np.asarray(x, np.float32)
    x = x - x.max(axis, keepdims=True)
    e = np.exp(x)
    return e / e.sum(axis, keepdims=True)


def top_p_distribution(logits, top_p: float, temperature: float) -> np.ndarray:
    """The categorical distribution top_p_sampling (sample_utils.py:3-39) draws from, as a dense
    [B, V] probability table: softmax(logits/T) -> sort desc -> keep tokens whose INCLUSIVE
    cumulative mass <= top_p -> renormalise.  If nothing is kept (p_1 > top_p; the reference then
    divides 0/0) the build's documented deviation applies: all mass on the argmax."""
    p = softmax(np.asarray(logits, np.float32) / np.float32(temperature))
    B, V = p.shape
    out = np.zeros_like(p)
    for b in range(B):
        order = np.argsort(-p[b], kind="stable")
        sp = p[b, order]
        cum = np.cumsum(sp, dtype=np.float32)
        keep = cum <= np.float32(top_p)
        if not keep.any():
            out[b, order[0]] = 1.0
            continue
        kept = np.where(keep, sp, 0)
        out[b, order] = kept / kept.sum()
    return out
