"""Kernel-level parity: every C-ABI entry point against the oracle on the same seeded inputs."""
import os

import numpy as np
import pytest
import torch

from oracle import mlx_ops as ref
from helpers import rel_err

pytestmark = pytest.mark.gpu

TD = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def t_(a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(TD[dtype]).cuda()


def w_(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).cuda()


def bits_of(t, dtype):
    if dtype == "float32":
        return t.cpu().numpy().view(np.uint32)
    return t.cpu().view(torch.int16).numpy().view(np.uint16)


@pytest.fixture(scope="module")
def ops():
    from mlx_parallm_b200 import ops as o, _lib
    _lib.require_device()
    return o


# ---------------------------------------------------------------- dequantize: bit-exact, all modes
@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("dtype", ["float32", "float16", "bfloat16"])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_dequantize_bit_exact_golden(ops, golden_dir, bits, dtype, mode):
    g = np.load(os.path.join(golden_dir, "dequantize_bits.npz"))
    out = ops.dequantize(w_(g[f"w_{bits}"]), t_(g[f"s_{bits}_{dtype}"], dtype), t_(g[f"b_{bits}_{dtype}"], dtype), 64, bits, mode)
    assert np.array_equal(bits_of(out, dtype), g[f"out_{bits}_{dtype}_{mode}"])


@pytest.mark.parametrize("bits,dtype", [(4, "float16"), (8, "bfloat16"), (4, "float32")])
def test_dequantize_bit_exact_random_large(ops, bits, dtype):
    rng = np.random.default_rng(11)
    N, K = 257, 4096
    wq = rng.integers(0, 2 ** 32, (N, K * bits // 32), dtype=np.uint32)
    s = ref.round_to(rng.uniform(-0.02, 0.02, (N, K // 64)), dtype)
    b = ref.round_to(rng.uniform(-0.2, 0.2, (N, K // 64)), dtype)
    for mode in (0, 1, 2):
        out = ops.dequantize(w_(wq), t_(s, dtype), t_(b, dtype), 64, bits, mode)
        exp = ref.storage_bits(ref.dequantize(wq, s, b, 64, bits, dtype, mode), dtype)
        assert np.array_equal(bits_of(out, dtype), exp), mode


def test_dequantize_empty_and_embed_gather(ops):
    rng = np.random.default_rng(12)
    V, K = 300, 128
    wq = rng.integers(0, 2 ** 32, (V, K // 8), dtype=np.uint32)
    s = ref.round_to(rng.uniform(-0.02, 0.02, (V, 2)), "float16")
    b = ref.round_to(rng.uniform(-0.2, 0.2, (V, 2)), "float16")
    out = ops.dequantize(w_(wq[:0]), t_(s[:0], "float16"), t_(b[:0], "float16"))
    assert out.shape == (0, K)
    ids = np.array([5, 0, 299, 5, 17], np.int32)
    e = ops.embed(torch.from_numpy(ids).cuda(), w_(wq), t_(s, "float16"), t_(b, "float16"))
    exp = ref.dequantize(wq[ids], s[ids], b[ids], 64, 4, "float16")
    assert e.dtype == torch.float32 and np.array_equal(e.cpu().numpy(), exp)


# ---------------------------------------------------------------- quantize: bit-exact (codes, scales, biases)
@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("dtype", ["float32", "float16", "bfloat16"])
@pytest.mark.parametrize("gs", [32, 64, 128])
def test_quantize_bit_exact(ops, bits, dtype, gs):
    rng = np.random.default_rng(bits * 100 + gs)
    N, K = 37, 512
    w = rng.standard_normal((N, K)).astype(np.float32) * 0.05
    w[0] = 0.0                                     # all-zero groups: scale falls back to eps, codes 0
    if dtype != "float16":
        w[1] = 0.031                               # constant groups (in f16 edge / eps overflows: NaN codes in MLX and the oracle alike)
    w[2, :gs] = -np.abs(w[2, :gs])                 # min side dominates -> positive scale
    w[3, :gs] = np.abs(w[3, :gs])                  # max side dominates -> negative scale
    w[4, ::2] = 0.0                                # zeros must be exactly representable
    w = ref.round_to(w, dtype)
    eq, es, eb = ref.quantize(w, gs, bits, dtype)
    gq, gsx, gb = ops.quantize(t_(w, dtype), gs, bits)
    assert np.array_equal(gq.cpu().numpy().view(np.uint32), eq)
    assert np.array_equal(bits_of(gsx, dtype), ref.storage_bits(es, dtype))
    assert np.array_equal(bits_of(gb, dtype), ref.storage_bits(eb, dtype))


def test_quantize_dequantize_roundtrip_full_size(ops):
    """Size-independent property at a BASELINE shape (down_proj of Llama-3-8B): dequantize(quantize(w)) is within half a
    quantisation step (+ storage rounding) of w, group by group, and quantising the dequantised matrix again reproduces
    the codes (idempotence)."""
    g = torch.Generator(device="cuda").manual_seed(4)
    w = (torch.randn((4096, 14336), generator=g, device="cuda") * 0.02).to(torch.float16)
    wq, s, b = ops.quantize(w, 64, 4)
    d = ops.dequantize(wq, s, b, 64, 4)
    step = s.float().abs().repeat_interleave(64, 1)
    # the scale is snapped to edge / rint(edge / scale), so the far end of a group may clip by up to ~0.8 step
    assert bool(((d.float() - w.float()).abs() <= 1.0 * step + 4e-3 * w.float().abs() + 1e-5).all())
    assert float(((d.float() - w.float()).abs() / step).mean()) < 0.3
    wq2, s2, b2 = ops.quantize(d, 64, 4)
    d2 = ops.dequantize(wq2, s2, b2, 64, 4)
    assert float((d2.float() - d.float()).abs().max()) <= float(step.max()) * 1.01


# ---------------------------------------------------------------- quantized matmul
def qmm_case(ops, M, N, K, bits, dtype, epi, seed=0, path=0, ragged_ld=False, small_scale=False):
    from mlx_parallm_b200 import _lib
    rng = np.random.default_rng(seed)
    wq = rng.integers(0, 2 ** 32, (N, K * bits // 32), dtype=np.uint32)
    nb = (1 << bits) - 1
    s = ref.round_to(rng.uniform(0.5, 1.5, (N, K // 64)) * rng.choice([-1, 1], (N, K // 64)) * (2 / nb) / np.sqrt(K), dtype)
    b = ref.round_to(-s * nb / 2 + rng.normal(0, 0.01 / np.sqrt(K), (N, K // 64)), dtype)
    x = ref.round_to(rng.standard_normal((M, K)), dtype)
    acc = ref.quantized_matmul_grouped(x, wq, s, b, 64, bits) if M * N * K <= 2 ** 24 else \
        ref.quantized_matmul(x, wq, s, b, 64, bits).astype(np.float64)
    xt = t_(x, dtype)
    if ragged_ld:
        big = torch.zeros((M, K + 64), dtype=xt.dtype, device="cuda")
        big[:, :K] = xt
        xt = big[:, :K]
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    _lib.check(_lib.lib().b200p_qmm_set_path(path))
    try:
        if epi == _lib.EPI_ADD_RESIDUAL:
            res = rng.standard_normal((M, N)).astype(np.float32)
            out = torch.from_numpy(res.copy()).cuda()
            ops.qmm(xt, w_(wq), t_(s, dtype), t_(b, dtype), 64, bits, epi, out=out, workspace=ws, small_scale=small_scale)
            exp = res + acc
        else:
            out = ops.qmm(xt, w_(wq), t_(s, dtype), t_(b, dtype), 64, bits, epi, workspace=ws, small_scale=small_scale)
            if epi == _lib.EPI_SILU_MUL:
                g, u = acc[:, :N // 2], acc[:, N // 2:]
                exp = g / (1 + np.exp(-g)) * u
            else:
                exp = acc
    finally:
        _lib.lib().b200p_qmm_set_path(0)
    return out.float().cpu().numpy(), exp


@pytest.mark.parametrize("dtype", ["float32", "float16", "bfloat16"])
@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("M", [1, 3, 8])
def test_qmv_store_matches_oracle(ops, dtype, bits, M):
    got, exp = qmm_case(ops, M, 96, 512, bits, dtype, 0, seed=M, path=1)
    tol = {"float32": 2e-5, "float16": 2e-3, "bfloat16": 1e-2}[dtype]     # output rounding of the storage type
    assert rel_err(got, exp) < tol


@pytest.mark.parametrize("epi", [1, 2, 3])
def test_qmv_epilogues(ops, epi):
    got, exp = qmm_case(ops, 5, 128, 256, 4, "float16", epi, seed=epi, path=1, ragged_ld=True)
    assert rel_err(got, exp) < (2e-3 if epi == 3 else 2e-5)


def test_qmv_tiny_model_shapes(ops):
    # K = 64 (one group per row, tiny hidden) and K = 128 (tiny down_proj), float32 like config 1
    for K in (64, 128):
        got, exp = qmm_case(ops, 8, 64, K, 4, "float32", 1, seed=K, path=1)
        assert rel_err(got, exp) < 2e-5


def test_qmm_rejects_bad_arguments(ops):
    from mlx_parallm_b200 import _lib
    x = torch.zeros((1, 96), dtype=torch.float16, device="cuda")
    w = torch.zeros((4, 12), dtype=torch.int32, device="cuda")
    s = torch.zeros((4, 1), dtype=torch.float16, device="cuda")
    with pytest.raises(_lib.B200PError):
        ops.qmm(x, w, s, s, 64, 4)                      # K=96 not a multiple of group 64


# ---------------------------------------------------------------- rmsnorm / rope+append / attention
@pytest.mark.parametrize("dtype", ["float32", "float16", "bfloat16"])
def test_rmsnorm(ops, dtype):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((7, 512)).astype(np.float32) * 3
    w = ref.round_to(1 + 0.1 * rng.standard_normal(512), dtype)
    y = ops.rmsnorm(torch.from_numpy(x).cuda(), t_(w, dtype), 1e-5)
    exp = ref.rms_norm(x, w, 1e-5, dtype)
    # identical up to 1 ulp of the storage type (fp32 reduction order differs)
    assert rel_err(y.float().cpu().numpy(), exp) < {"float32": 1e-6, "float16": 1.1e-3, "bfloat16": 8.1e-3}[dtype]


@pytest.mark.parametrize("dtype,M,H", [
    ("float16", 8, 4096),       # one-pass cluster kernel at a few rows
    ("float16", 1, 4096),
    ("bfloat16", 8, 5120),      # below 16 rows and no cluster form: the 256-thread kernel
    ("float16", 64, 4096),      # early one-pass kernel, cluster of 2 (Llama-3-8B decode at B=64)
    ("bfloat16", 32, 8192),     # cluster of 4 (70B-class)
    ("float16", 16, 2048),      # cluster of 1
    ("bfloat16", 64, 5120),     # Qwen3-14B: 5120 is not a multiple of 2048 -> early two-pass kernel
    ("float16", 100, 4096),     # 100 rows x 2 CTAs exceed the SMs -> early two-pass kernel
])
@pytest.mark.parametrize("perm", [False, True])
def test_rmsnorm_variants_at_decode_sizes(ops, dtype, M, H, perm):
    """Every RMSNorm variant the dispatcher can pick (csrc/elementwise.cu), PERM8 and plain, against the oracle."""
    rng = np.random.default_rng(31)
    x = rng.standard_normal((M, H)).astype(np.float32) * 3
    x[M // 2] *= 50.0                               # rows of very different norms
    w = ref.round_to(1 + 0.1 * rng.standard_normal(H), dtype)
    y = ops.rmsnorm(torch.from_numpy(x).cuda(), t_(w, dtype), 1e-5, out_perm=perm).float().cpu().numpy()
    if perm:                                        # undo [0,4,1,5,2,6,3,7] inside every aligned group of 8
        y = y.reshape(M, H // 8, 8)[:, :, [0, 2, 4, 6, 1, 3, 5, 7]].reshape(M, H)
    exp = ref.rms_norm(x, w, 1e-5, dtype)
    for r in range(M):                              # per row: a row of small norm must not hide behind a large one
        assert rel_err(y[r], exp[r]) < {"float16": 1.1e-3, "bfloat16": 8.1e-3}[dtype], r


@pytest.mark.parametrize("dtype,D,qk_norm", [("float16", 128, False), ("bfloat16", 128, True), ("float32", 16, False), ("float16", 64, True)])
def test_rope_kv_append(ops, dtype, D, qk_norm):
    rng = np.random.default_rng(4)
    B, L, Hq, Hkv, Tcap = 3, 2, 4, 2, 40
    lengths = np.array([0, 7, 33], np.int32)
    qkv = ref.round_to(rng.standard_normal((B * L, (Hq + 2 * Hkv) * D)), dtype)
    qn = ref.round_to(1 + 0.1 * rng.standard_normal(D), dtype)
    kn = ref.round_to(1 + 0.1 * rng.standard_normal(D), dtype)
    q_out = torch.zeros((B, L, Hq, D), dtype=TD[dtype], device="cuda")
    kc = torch.zeros((B, Hkv, Tcap, D), dtype=TD[dtype], device="cuda")
    vc = torch.zeros_like(kc)
    ops.rope_kv_append(t_(qkv, dtype), q_out, kc, vc, torch.from_numpy(lengths).cuda(), B, L, Hq, Hkv, D, 10000.0, 1.0,
                       t_(qn, dtype) if qk_norm else None, t_(kn, dtype) if qk_norm else None, 1e-6)
    r = qkv.reshape(B, L, Hq + 2 * Hkv, D)
    q, k, v = r[:, :, :Hq], r[:, :, Hq:Hq + Hkv], r[:, :, Hq + Hkv:]
    if qk_norm:
        q = ref.rms_norm(q, qn, 1e-6, dtype)
        k = ref.rms_norm(k, kn, 1e-6, dtype)
    qe = ref.rope(q.transpose(0, 2, 1, 3), lengths, 10000.0, 1.0, dtype).transpose(0, 2, 1, 3)
    ke = ref.rope(k.transpose(0, 2, 1, 3), lengths, 10000.0, 1.0, dtype)
    tol = {"float32": 3e-6, "float16": 1.5e-3, "bfloat16": 1e-2}[dtype]
    assert rel_err(q_out.float().cpu().numpy(), qe) < tol
    kcn, vcn = kc.float().cpu().numpy(), vc.float().cpu().numpy()
    for b in range(B):
        s = lengths[b]
        assert rel_err(kcn[b, :, s:s + L], ke[b]) < tol
        assert np.array_equal(vcn[b, :, s:s + L], v[b].transpose(1, 0, 2))
        assert not kcn[b, :, s + L:].any() and not kcn[b, :, :s].any()      # nothing else touched


@pytest.mark.parametrize("D", [128, 64])          # the head_dim-128 warp kernel and the generic one
def test_rope_kv_append_counts_rows_beyond_capacity(ops, D):
    """A row whose position reached Tcap cannot be appended: the kernel drops it and bumps the per-device counter that
    b200p_kv_overflow_count reads (and the generation loop checks); rows with room are written as usual."""
    B, L, Hq, Hkv, Tcap = 3, 1, 4, 2, 8
    ops.kv_overflow_count(reset=True)
    lengths = torch.tensor([3, 8, 9], dtype=torch.int32, device="cuda")       # row 0 fits, rows 1 and 2 are full
    qkv = torch.randn((B * L, (Hq + 2 * Hkv) * D), device="cuda").to(torch.float16)
    q_out = torch.zeros((B, L, Hq, D), dtype=torch.float16, device="cuda")
    kc = torch.zeros((B, Hkv, Tcap, D), dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    ops.rope_kv_append(qkv, q_out, kc, vc, lengths, B, L, Hq, Hkv, D, 10000.0)
    assert ops.kv_overflow_count() == 2 * Hkv                 # one count per dropped (token, V head)
    assert ops.kv_overflow_count(reset=True) == 2 * Hkv       # reading does not clear; reset returns the old value
    assert ops.kv_overflow_count() == 0
    v_row0 = qkv[0].view(Hq + 2 * Hkv, D)[Hq + Hkv:]
    assert torch.equal(vc[0, :, 3], v_row0) and not vc[1:].any() and not kc[1:].any()
    ops.rope_kv_append(qkv[:1], q_out[:1], kc[:1], vc[:1], lengths[:1], 1, L, Hq, Hkv, D, 10000.0)
    assert ops.kv_overflow_count() == 0                       # nothing dropped: the counter stays clear


def attn_case(ops, B, L, Hq, Hkv, D, lengths, dtype, path, seed=5):
    from mlx_parallm_b200 import _lib
    rng = np.random.default_rng(seed)
    Tcap = int(max(lengths)) + L + 3
    q = ref.round_to(rng.standard_normal((B, L, Hq, D)), dtype)
    k = ref.round_to(rng.standard_normal((B, Hkv, Tcap, D)), dtype)
    v = ref.round_to(rng.standard_normal((B, Hkv, Tcap, D)), dtype)
    out = torch.zeros((B, L, Hq * D), dtype=TD[dtype], device="cuda")
    ws = torch.empty((max(ops.attn_workspace_bytes(B, L, Hq, D), 16),), dtype=torch.uint8, device="cuda")
    _lib.check(_lib.lib().b200p_attn_set_path(path))
    try:
        ops.attn(t_(q, dtype), t_(k, dtype), t_(v, dtype), torch.from_numpy(np.asarray(lengths, np.int32)).cuda(), out,
                 B, L, Hq, Hkv, D, D ** -0.5, workspace=ws)
    finally:
        _lib.lib().b200p_attn_set_path(0)
    exp = np.zeros((B, L, Hq * D), np.float32)
    for b in range(B):
        T = lengths[b] + L
        mask = ref.additive_causal_mask_variable(L, [lengths[b]], T)
        o = ref.sdpa(q[b:b + 1].transpose(0, 2, 1, 3), k[b:b + 1, :, :T], v[b:b + 1, :, :T], D ** -0.5, mask if L > 1 else None)
        exp[b] = o.transpose(0, 2, 1, 3).reshape(L, Hq * D)
    return out.float().cpu().numpy(), exp


@pytest.mark.parametrize("dtype,D,Hq,Hkv", [("float32", 16, 4, 4), ("float16", 128, 8, 2), ("bfloat16", 128, 5, 1), ("float16", 64, 4, 2)])
def test_attention_generic_decode_ragged_lengths(ops, dtype, D, Hq, Hkv):
    got, exp = attn_case(ops, 3, 1, Hq, Hkv, D, [0, 5, 130], dtype, path=1)
    assert rel_err(got, exp) < {"float32": 2e-5, "float16": 1.5e-3, "bfloat16": 1e-2}[dtype]


def test_attention_generic_prefill_causal(ops):
    got, exp = attn_case(ops, 2, 9, 4, 2, 64, [0, 4], "float16", path=1)
    assert rel_err(got, exp) < 1.5e-3


@pytest.mark.parametrize("dtype,B,L,Hq,Hkv,lengths", [
    ("float16", 2, 70, 8, 2, [0, 37]),            # two query tiles per row, ragged cache offsets
    ("bfloat16", 3, 9, 5, 1, [0, 4, 130]),        # fewer queries than a tile, G = 5
    ("float16", 1, 200, 4, 4, [64]),              # several key tiles, MHA
    ("float16", 2, 64, 8, 8, [0, 0]),             # exactly one query tile, pure causal
])
def test_attention_tensor_prefill_causal(ops, dtype, B, L, Hq, Hkv, lengths):
    """mma.sync flash prefill (attn_prefill.cu) against the oracle's SDPA with the additive causal mask."""
    got, exp = attn_case(ops, B, L, Hq, Hkv, 128, lengths, dtype, path=2)
    assert rel_err(got, exp) < {"float16": 2e-3, "bfloat16": 1.2e-2}[dtype]
    gen, _ = attn_case(ops, B, L, Hq, Hkv, 128, lengths, dtype, path=1)
    assert rel_err(got, gen) < {"float16": 2e-3, "bfloat16": 1.2e-2}[dtype]


# ---------------------------------------------------------------- samplers
def test_argmax_ties_and_probs(ops):
    rng = np.random.default_rng(6)
    lg = rng.standard_normal((5, 1000)).astype(np.float32)
    lg[0, 17] = lg[0, 900] = 9.0                       # tie -> lowest index
    lg[1, 999] = 12.0
    probs = torch.zeros(5, device="cuda")
    tok = ops.argmax(torch.from_numpy(lg).cuda(), probs_out=probs)
    assert tok.tolist() == ref.argmax_tokens(lg)[:, 0].tolist() and tok[0].item() == 17
    sm = ref.softmax(lg)
    assert np.allclose(probs.cpu().numpy(), sm[np.arange(5), tok.cpu().numpy()], rtol=1e-4)


def test_argmax_large_vocab_unaligned(ops):
    rng = np.random.default_rng(7)
    lg = rng.standard_normal((3, 151937)).astype(np.float32)
    t = torch.from_numpy(lg).cuda()
    assert ops.argmax(t).tolist() == lg.argmax(-1).tolist()


@pytest.mark.parametrize("top_p,temp", [(0.9, 1.0), (0.5, 0.7), (1.0, 1.3)])
def test_top_p_sampling_distribution(ops, top_p, temp):
    """Distributional parity (A.7: MLX's threefry stream is not reproducible outside MLX): draws must stay in
    the oracle's nucleus and match its renormalised probabilities (chi-square style bound on the big bins)."""
    rng = np.random.default_rng(8)
    V, n = 200, 20000
    lg = (rng.standard_normal(V) * 2.0).astype(np.float32)
    dist = ref.top_p_distribution(lg[None], top_p, temp)[0]
    logits = torch.from_numpy(np.tile(lg, (n, 1))).cuda()
    tok = ops.sample_top_p(logits, temp, top_p, seed=123, offset=0).cpu().numpy()
    assert dist[tok].min() > 0, "sampled a token outside the nucleus"
    freq = np.bincount(tok, minlength=V) / n
    big = dist > 0.01
    assert np.abs(freq[big] - dist[big]).max() < 4 * np.sqrt(dist[big].max() / n) + 2e-3
    tok2 = ops.sample_top_p(logits, temp, top_p, seed=123, offset=0).cpu().numpy()
    assert np.array_equal(tok, tok2)                     # counter-based RNG: reproducible
    tok3 = ops.sample_top_p(logits, temp, top_p, seed=123, offset=1).cpu().numpy()
    assert not np.array_equal(tok, tok3)


@pytest.mark.parametrize("V,top_p,temp", [(128256, 0.9, 0.7), (151937, 0.6, 1.0), (32000, 0.95, 1.2)])
def test_top_p_large_vocab_stays_in_the_oracle_nucleus(ops, V, top_p, temp):
    """BASELINE vocab sizes (151 937 = unaligned rows -> scalar loads): LM-like logits (a gaussian body, a few large values,
    a block of exactly equal logits as untrained vocab slots have), 1500 draws: every token inside the oracle's nucleus,
    frequencies of the heavy tokens as the oracle's renormalised distribution says."""
    rng = np.random.default_rng(V)
    n = 1500
    lg = (rng.standard_normal(V) * 2.0).astype(np.float32)
    lg[rng.choice(V, 40, replace=False)] = (8.0 + 1.5 * rng.random(40)).astype(np.float32)
    lg[1000:3000] = -3.25
    dist = ref.top_p_distribution(lg[None], top_p, temp)[0]
    assert 5 < (dist > 0).sum() < V
    logits = torch.from_numpy(lg).cuda().repeat(n, 1)
    probs = torch.zeros(n, device="cuda")
    tok = ops.sample_top_p(logits, temp, top_p, seed=5, offset=9, probs_out=probs).cpu().numpy()
    # the nucleus ends among thousands of tokens of ~1e-5 each: WHICH token is the last one kept depends on the summation
    # order of ~1e4 float32 terms (numpy's sequential cumsum vs the kernel's tree), so membership is checked against the
    # float64 cumulative mass with a 1e-4 band at the boundary, the frequencies against the oracle's table
    p64 = np.exp((lg.astype(np.float64) - lg.max()) / temp)
    p64 /= p64.sum()
    order = np.argsort(-p64, kind="stable")
    cum = np.empty(V)
    cum[order] = np.cumsum(p64[order])
    assert cum[tok].max() <= top_p + 1e-4, "sampled a token outside the nucleus"
    freq = np.bincount(tok, minlength=V) / n
    big = dist > 0.008
    assert big.sum() >= 3 and np.abs(freq[big] - dist[big]).max() < 4 * np.sqrt(dist[big].max() / n) + 5e-3
    sm = ref.softmax(lg[None])[0]
    assert np.allclose(probs.cpu().numpy(), sm[tok], rtol=2e-3)        # probs = softmax(logits)[token] (utils.py:363)
    # top_p = 1: plain categorical over the whole vocabulary
    tok1 = ops.sample_top_p(logits, temp, 1.0, seed=5, offset=9).cpu().numpy()
    d1 = ref.top_p_distribution(lg[None], 1.0, temp)[0]
    big1 = d1 > 0.008
    f1 = np.bincount(tok1, minlength=V) / n
    assert np.abs(f1[big1] - d1[big1]).max() < 4 * np.sqrt(d1[big1].max() / n) + 5e-3


@pytest.mark.parametrize("B", [1, 30, 70, 200])
def test_samplers_at_every_cluster_size(ops, B):
    """Rows are split over thread-block clusters of 8 / 4 / 2 / 1 CTAs depending on the batch (sampler.cu cluster_for): same
    argmax (ties -> lowest index, also across segment borders), same softmax prob, draws inside the nucleus at each size."""
    rng = np.random.default_rng(B)
    V = 40003                                            # unaligned rows: scalar loads, ragged last segment
    lg = (rng.standard_normal((B, V)) * 2.0).astype(np.float32)
    lg[:, 77] = lg[:, 39000] = 11.0                      # tie between the first and the last segment
    if B > 1:
        lg[1, 39000] = 12.0
    t = torch.from_numpy(lg).cuda()
    probs = torch.zeros(B, device="cuda")
    tok = ops.argmax(t, probs_out=probs).cpu().numpy()
    assert tok.tolist() == lg.argmax(-1).tolist() and tok[0] == 77
    sm = ref.softmax(lg)
    assert np.allclose(probs.cpu().numpy(), sm[np.arange(B), tok], rtol=1e-4)
    lg[:, rng.choice(V, 30, replace=False)] = (7.0 + 2.0 * rng.random(30)).astype(np.float32)
    t = torch.from_numpy(lg).cuda()
    for trial in range(3):
        tk = ops.sample_top_p(t, 0.8, 0.9, seed=11, offset=trial, probs_out=probs).cpu().numpy()
        p64 = np.exp((lg.astype(np.float64) - lg.max(-1, keepdims=True)) / 0.8)
        p64 /= p64.sum(-1, keepdims=True)
        for b in range(B):
            order = np.argsort(-p64[b], kind="stable")
            cum = np.cumsum(p64[b, order])
            rank = int(np.nonzero(order == tk[b])[0][0])
            assert cum[rank] <= 0.9 + 1e-4, (b, tk[b])
        assert np.allclose(probs.cpu().numpy(), ref.softmax(lg)[np.arange(B), tk], rtol=2e-3)


def test_top_p_exact_boundary_and_ties(ops):
    """cum <= top_p keeps exactly the prefix the reference keeps (sample_utils.py:25-31): p = .4 .3 .2 .1 with top_p .75 ->
    {0, 1}; top_p .95 -> {0, 1, 2}.  A nucleus that ends inside a run of EQUAL probabilities: the reference keeps as many of the run as fit, in
    whatever order its argsort leaves ties; this kernel selects by value and drops the run (documented in sampler.cu)."""
    p = np.array([0.4, 0.3, 0.2, 0.1], np.float32)
    lg = torch.from_numpy(np.tile(np.log(p), (4096, 1))).cuda()
    t = ops.sample_top_p(lg, 1.0, 0.75, 3, 0).cpu().numpy()
    assert set(t.tolist()) == {0, 1} and abs((t == 0).mean() - 4 / 7) < 0.04
    t = ops.sample_top_p(lg, 1.0, 0.95, 3, 1).cpu().numpy()
    assert set(t.tolist()) == {0, 1, 2}
    q = np.array([0.4, 0.2, 0.2, 0.2], np.float32)          # cum: .4 .6 .8 1.0, top_p .7 cuts inside the run of .2s
    lq = torch.from_numpy(np.tile(np.log(q), (1024, 1))).cuda()
    t = ops.sample_top_p(lq, 1.0, 0.7, 3, 2).cpu().numpy()
    assert set(t.tolist()) <= {0, 1, 2, 3} and (t == 0).mean() > 0.6


def test_top_p_empty_nucleus_returns_argmax(ops):
    lg = np.log(np.array([[0.6, 0.3, 0.1]], np.float32))
    tok = ops.sample_top_p(torch.from_numpy(np.tile(lg, (64, 1))).cuda(), 1.0, 0.5, 1, 0)
    assert (tok.cpu().numpy() == 0).all()


# ---------------------------------------------------------------- tcgen05 tile path of b200p_qmm
TC_CASES = [  # M, N, K, bits, dtype, epilogue
    (8, 256, 256, 4, "float16", 1), (16, 512, 1024, 4, "float16", 0), (64, 1024, 512, 4, "float16", 2),
    (64, 2048, 1024, 4, "float16", 3), (5, 1000, 192, 4, "float16", 1), (100, 1000, 448, 4, "bfloat16", 1),
    (300, 512, 512, 4, "float16", 0), (64, 512, 2048, 8, "float16", 1), (33, 768, 640, 4, "bfloat16", 3),
    (64, 256, 4096, 4, "float16", 2),          # deep K with few row blocks -> split-K + last-arriver fix-up
    # 8-bit bf16: the code goes in as two nibbles (128 + q is exact in bf16 only below 128), two bf16 FMAs
    (64, 512, 2048, 8, "bfloat16", 1), (8, 1024, 1024, 8, "bfloat16", 2), (40, 768, 512, 8, "bfloat16", 3), (128, 512, 1024, 8, "bfloat16", 0),
]


@pytest.mark.parametrize("M,N,K,bits,dtype,epi", TC_CASES)
def test_qmm_tensor_path_matches_oracle(ops, M, N, K, bits, dtype, epi):
    got, exp = qmm_case(ops, M, N, K, bits, dtype, epi, seed=M + N, path=2)
    # A is the storage-rounded dequantised weight (round_T(q*s+b)), accumulation fp32 in TMEM
    tol = {"float16": 3e-3, "bfloat16": 2e-2}[dtype]
    assert rel_err(got, exp) < tol


def test_qmm_tensor_path_is_bit_reproducible_and_matches_gemv(ops):
    # whole-tile kernel (STORE / STORE_F32 / SILU_MUL): K splits are reduced from per-CTA partial tiles in CTA order (no
    # atomics): two runs agree bit for bit.  The fp32 accumulate epilogue runs on the stream-K kernel, whose partial tiles
    # are summed by the L2 atomic units in arrival order: equal up to fp32 summation order only.
    for epi in (1, 3, 0):
        a1, _ = qmm_case(ops, 64, 512, 4096, 4, "float16", epi, seed=3, path=2)
        a2, _ = qmm_case(ops, 64, 512, 4096, 4, "float16", epi, seed=3, path=2)
        assert np.array_equal(a1, a2), epi
    a1, _ = qmm_case(ops, 64, 512, 4096, 4, "float16", 2, seed=3, path=2)
    a2, _ = qmm_case(ops, 64, 512, 4096, 4, "float16", 2, seed=3, path=2)
    assert rel_err(a1, a2) < 1e-6
    g, _ = qmm_case(ops, 64, 512, 4096, 4, "float16", 1, seed=3, path=1)
    a1, _ = qmm_case(ops, 64, 512, 4096, 4, "float16", 1, seed=3, path=2)
    assert rel_err(a1, g) < 1e-3


SPLIT_CASES = [  # every CTA holds a head AND a tail partial tile, tiles shared by 2..8 CTAs, ragged token / row edges
    (64, 4096, 4096, 4, "float16", 2), (64, 6144, 4096, 4, "float16", 0), (64, 28672 // 4, 4096, 4, "float16", 3),
    (64, 4096, 14336 // 2, 4, "float16", 2), (37, 1000, 2048, 4, "bfloat16", 1), (128, 1024, 4096, 4, "float16", 3),
    (130, 520, 1536, 4, "float16", 0), (8, 4096, 4096, 8, "float16", 2), (1, 768, 8192, 4, "float16", 1),
]


@pytest.mark.parametrize("M,N,K,bits,dtype,epi", SPLIT_CASES)
def test_qmm_tensor_path_split_tiles(ops, M, N, K, bits, dtype, epi):
    """Layer-sized shapes where stream-K splits every tile: partial tiles through the workspace, shared final reduction,
    counters left re-armed (the second call reuses the same workspace inside qmm_case's fresh allocation, the third
    test below reuses one workspace across shapes)."""
    got, exp = qmm_case(ops, M, N, K, bits, dtype, epi, seed=M + N + K, path=2)
    tol = {"float16": 3e-3, "bfloat16": 2e-2}[dtype]
    assert rel_err(got, exp) < tol


def test_qmm_tensor_path_workspace_is_left_rearmed(ops):
    """One workspace, many launches of different shapes back to back (what a decode step does): the per-tile counters
    must be back at zero after every launch, stale partial tiles must never be read."""
    from mlx_parallm_b200 import _lib
    rng = np.random.default_rng(5)
    shapes = [(64, 4096, 4096, 2), (64, 6144, 4096, 0), (64, 2048, 1024, 3), (16, 512, 8192, 1), (64, 4096, 4096, 2)]
    ws = torch.zeros((max(ops.qmm_workspace_bytes(64, n, k) for _, n, k, _ in shapes),), dtype=torch.uint8, device="cuda")
    _lib.check(_lib.lib().b200p_qmm_set_path(2))
    try:
        for rep in range(3):
            for M, N, K, epi in shapes:
                wq = rng.integers(0, 2 ** 32, (N, K // 8), dtype=np.uint32)
                s = ref.round_to(rng.uniform(0.5, 1.5, (N, K // 64)) * (2 / 15) / np.sqrt(K), "float16")
                b = ref.round_to(-s * 7.5, "float16")
                x = ref.round_to(rng.standard_normal((M, K)), "float16")
                acc = ref.quantized_matmul(x, wq, s, b, 64, 4).astype(np.float64)
                if epi == 2:
                    out = torch.zeros((M, N), dtype=torch.float32, device="cuda")
                    ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, epi, out=out, workspace=ws)
                    exp = acc
                else:
                    out = ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, epi, workspace=ws)
                    exp = acc if epi != 3 else acc[:, :N // 2] / (1 + np.exp(-acc[:, :N // 2])) * acc[:, N // 2:]
                assert rel_err(out.float().cpu().numpy(), exp) < 3e-3, (rep, M, N, K, epi)
        torch.cuda.synchronize()
        ncnt = 2 * ((4096 + 255) // 256 + 1)
        assert int(ws[:ncnt * 4].view(torch.int32).abs().sum()) == 0
    finally:
        _lib.lib().b200p_qmm_set_path(0)


@pytest.mark.parametrize("M,N,K,bits,epi", [(64, 2048, 1024, 4, 3), (64, 4096, 4096, 4, 2), (8, 256, 256, 4, 1), (16, 512, 1024, 4, 0),
                                            (64, 512, 2048, 8, 1), (5, 1000, 192, 4, 1), (200, 512, 512, 4, 0)])
def test_qmm_small_scale_dequant_is_bit_identical(ops, M, N, K, bits, epi):
    """B200P_W_SMALLSCALE (one half2 FMA per code pair on subnormal operands, accumulator rescaled by 16) rounds q*s+b at
    the same point as the 3-op dequant: identical outputs wherever the operand / 16 stays a normal half."""
    a, exp = qmm_case(ops, M, N, K, bits, "float16", epi, seed=11, path=2, small_scale=False)
    f, _ = qmm_case(ops, M, N, K, bits, "float16", epi, seed=11, path=2, small_scale=True)
    assert rel_err(f, exp) < 3e-3
    if bits == 4 and K <= 1024:
        assert np.array_equal(a, f)          # every |q*s+b| / 16 is a normal half here
    else:
        assert rel_err(a, f) < 2e-5          # operands below 2^-10 round as subnormals at 1/16: absolute error <= 2^-21


# ---------------------------------------------------------------- mma.sync register GEMV of b200p_qmm (M <= 16)
MMA_CASES = [  # M, N, K, dtype, epilogue (1 = fp32 store, 2 = += into fp32, split-K with red.add)
    (1, 256, 256, "float16", 1), (3, 512, 1024, "float16", 2), (8, 96, 512, "bfloat16", 1), (9, 1024, 4096, "float16", 2),
    (16, 2048, 256, "bfloat16", 2), (5, 64, 768, "float16", 1), (2, 4096, 4096, "float16", 2), (13, 160, 1280, "float16", 1),
]


@pytest.mark.parametrize("M,N,K,dtype,epi", MMA_CASES)
def test_qmm_mma_gemv_matches_oracle(ops, M, N, K, dtype, epi):
    got, exp = qmm_case(ops, M, N, K, 4, dtype, epi, seed=M + N + K, path=3, ragged_ld=(M == 5))
    tol = {"float16": 3e-3, "bfloat16": 2e-2}[dtype]     # A = storage-rounded q*s+b, fp32 accumulation in the MMA
    assert rel_err(got, exp) < tol


def test_qmm_mma_gemv_agrees_with_the_other_paths_and_is_the_auto_choice(ops):
    from mlx_parallm_b200 import _lib
    m, _ = qmm_case(ops, 4, 512, 2048, 4, "float16", 1, seed=9, path=3)
    t, _ = qmm_case(ops, 4, 512, 2048, 4, "float16", 1, seed=9, path=2) if \
        _lib.lib().b200p_qmm_tensor_supported(4, 512, 2048, 64, 4, 1, 1) else (m, None)
    g, _ = qmm_case(ops, 4, 512, 2048, 4, "float16", 1, seed=9, path=1)
    a, _ = qmm_case(ops, 4, 512, 2048, 4, "float16", 1, seed=9, path=0)
    assert rel_err(m, t) < 1e-5 and rel_err(m, g) < 1e-3
    assert np.array_equal(a, m)                          # small matrix, M <= 4: auto == the mma.sync GEMV, bit for bit
    assert _lib.lib().b200p_qmm_set_path(5) == -1


@pytest.mark.parametrize("dtype,M", [("float16", 1), ("bfloat16", 3), ("float16", 8), ("float16", 13)])
def test_qmv_mma_silu_mul_pairs_gate_and_up_rows(ops, dtype, M):
    """SILU_MUL through the mma.sync GEMV (forced path 3): a warp takes 8 gate rows and the 8 matching up rows."""
    from mlx_parallm_b200 import _lib
    got, exp = qmm_case(ops, M, 256, 512, 4, dtype, _lib.EPI_SILU_MUL, seed=40 + M, path=3)
    assert got.shape == (M, 128)
    assert rel_err(got, exp) < {"float16": 2e-3, "bfloat16": 1e-2}[dtype]
    tile, _ = qmm_case(ops, M, 256, 512, 4, dtype, _lib.EPI_SILU_MUL, seed=40 + M, path=2)
    assert rel_err(got, tile) < {"float16": 1.1e-3, "bfloat16": 8.1e-3}[dtype]      # one rounding of the same fp32 value


# ---------------------------------------------------------------- mma.sync small-matrix GEMM (csrc/qmm_mma.cu)
@pytest.mark.parametrize("dtype,M,N,K", [
    ("float16", 5, 128, 512),        # NT = 1, two steps, no K split
    ("float16", 8, 256, 1024),
    ("bfloat16", 16, 128, 2048),     # NT = 2
    ("float16", 23, 384, 1024),      # NT = 4 with zero-filled token lines
    ("float16", 64, 256, 2048),      # NT = 8, K split
    ("bfloat16", 64, 128, 256),      # a single step
    ("float16", 33, 1024, 4096),     # many row blocks x splits: partial sums meet through the bulk reduce-add
])
def test_qmm_mma_small_matrix_gemm_matches_oracle(ops, dtype, M, N, K):
    from mlx_parallm_b200 import _lib
    got, exp = qmm_case(ops, M, N, K, 4, dtype, _lib.EPI_ADD_RESIDUAL, seed=M + N, path=4, ragged_ld=(M == 23))
    assert rel_err(got, exp) < {"float16": 1e-3, "bfloat16": 6e-3}[dtype]     # bound: rounding of the dequantised weights
    auto, _ = qmm_case(ops, M, N, K, 4, dtype, _lib.EPI_ADD_RESIDUAL, seed=M + N, path=0)
    assert rel_err(auto, got) < 1e-5                     # auto (this kernel up to 16 rows, stream-K tiles above): same dequantised
                                                         # weights, fp32 sums in another order


def test_qmm_mma_perm8_activations_and_unsupported_shapes(ops):
    from mlx_parallm_b200 import _lib
    from test_gpu_layouts import to_perm8
    rng = np.random.default_rng(77)
    M, N, K = 64, 256, 1024
    wq = rng.integers(0, 2 ** 32, (N, K // 8), dtype=np.uint32)
    s = ref.round_to(rng.uniform(0.5, 1.5, (N, K // 64)) * (2 / 15) / np.sqrt(K), "float16")
    b = ref.round_to(-s * 7.5, "float16")
    x = ref.round_to(rng.standard_normal((M, K)), "float16")
    _lib.check(_lib.lib().b200p_qmm_set_path(4))
    try:
        nat = torch.zeros((M, N), dtype=torch.float32, device="cuda")
        per = torch.zeros((M, N), dtype=torch.float32, device="cuda")
        ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, _lib.EPI_ADD_RESIDUAL, out=nat)
        ops.qmm(t_(to_perm8(x), "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, _lib.EPI_ADD_RESIDUAL, out=per,
                x_perm=True)
        assert rel_err(per.cpu().numpy(), nat.cpu().numpy()) < 1e-6
        assert rel_err(nat.cpu().numpy(), ref.quantized_matmul(x, wq, s, b, 64, 4)) < 1e-3
        with pytest.raises(_lib.B200PError):                  # N not a multiple of 128: forced path refuses, nothing is launched
            ops.qmm(t_(x, "float16"), w_(wq[:96]), t_(s[:96], "float16"), t_(b[:96], "float16"), 64, 4, _lib.EPI_ADD_RESIDUAL,
                    out=nat[:, :96].contiguous())
        with pytest.raises(_lib.B200PError):                  # STORE epilogue: not this kernel's
            ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, _lib.EPI_STORE_F32)
    finally:
        _lib.lib().b200p_qmm_set_path(0)


# ---------------------------------------------------------------- split-K tensor decode attention
@pytest.mark.parametrize("dtype,Hq,Hkv,lengths", [
    ("float16", 8, 2, [0, 5, 130]),             # ragged rows incl. an empty cache (only the new token)
    ("bfloat16", 5, 1, [15, 16, 17]),           # G = 5 (Qwen3-14B at TP=8), tile-boundary lengths
    ("float16", 32, 8, [700, 3]),               # Llama-3-8B head layout
    ("float16", 8, 1, [5000, 4999, 1, 2500]),   # G = 8 (70B), long rows -> sequence split across CTAs
])
def test_attention_tensor_decode(ops, dtype, Hq, Hkv, lengths):
    got, exp = attn_case(ops, len(lengths), 1, Hq, Hkv, 128, lengths, dtype, path=2)
    assert rel_err(got, exp) < {"float16": 2e-3, "bfloat16": 1.2e-2}[dtype]


def test_attention_tensor_matches_generic(ops):
    a, _ = attn_case(ops, 2, 1, 8, 2, 128, [1000, 37], "float16", path=2, seed=9)
    g, _ = attn_case(ops, 2, 1, 8, 2, 128, [1000, 37], "float16", path=1, seed=9)
    assert rel_err(a, g) < 2e-3
