"""Known-answer tests that pin the oracle: hand-derived, independent of the oracle's own code path."""
import numpy as np
import pytest

from oracle import mlx_ops as ops
from oracle.model import OracleConfig, OracleModel, OraclePagedKVCache, generate_step
from helpers import make_pair, prompts_for, rel_err


# ---- packing (A.1) ---------------------------------------------------------------------------------
def test_pack_layout_4bit_little_end_first():
    q = np.array([[1, 2, 3, 4, 5, 6, 7, 8]], dtype=np.uint32)
    assert ops.pack_codes(q, 4)[0, 0] == 0x87654321
    assert (ops.unpack_codes(np.array([[0x87654321]], np.uint32), 4) == q).all()


def test_pack_layout_8bit():
    q = np.array([[0x11, 0x22, 0x33, 0xFF]], dtype=np.uint32)
    assert ops.pack_codes(q, 8)[0, 0] == 0xFF332211


def test_pack_roundtrip_random():
    rng = np.random.default_rng(0)
    for bits in (4, 8):
        q = rng.integers(0, 1 << bits, (5, 128)).astype(np.uint32)
        assert (ops.unpack_codes(ops.pack_codes(q, bits), bits) == q).all()


# ---- quantize (A.1) ---------------------------------------------------------------------------------
def test_quantize_positive_edge_branch():
    # group = linspace(-1, 2): |min| < |max| -> scale negative, edge = max = 2; scale0 = -(3/15) = -0.2;
    # q0 = rint(2 / -0.2) = -10 -> scale = 2 / -10 = -0.2, bias = 2.  w=2 -> q=0; w=-1 -> q=15.
    w = np.linspace(-1, 2, 64, dtype=np.float32)[None]
    wq, s, b = ops.quantize(w, 64, 4, "float32")
    assert s[0, 0] == pytest.approx(-0.2, rel=1e-6) and b[0, 0] == pytest.approx(2.0)
    q = ops.unpack_codes(wq, 4)[0]
    assert q[-1] == 0 and q[0] == 15


def test_quantize_negative_edge_branch_and_zero_exact():
    # group = linspace(-2, 1): |min| > |max| -> scale positive 0.2, edge = min = -2, q0 = -10 -> bias -2.
    w = np.linspace(-2, 1, 64, dtype=np.float32)[None]
    wq, s, b = ops.quantize(w, 64, 4, "float32")
    assert s[0, 0] == pytest.approx(0.2, rel=1e-6) and b[0, 0] == pytest.approx(-2.0)
    # zero is exactly representable: code 10 -> 10*0.2 - 2 == 0 up to fp32 rounding of 0.2
    d = ops.dequantize(wq, s, b, 64, 4, "float32")
    assert np.abs(d).min() < 1e-6
    # quantisation error bounded by half a step
    assert np.abs(d - w).max() <= 0.1 + 1e-6


def test_quantize_all_zero_group():
    w = np.zeros((1, 64), np.float32)
    wq, s, b = ops.quantize(w, 64, 4, "float32")
    assert np.all(ops.dequantize(wq, s, b, 64, 4, "float32") == 0)


# ---- dequantize (A.2) -------------------------------------------------------------------------------
def test_dequantize_known_values_and_modes():
    wq = ops.pack_codes(np.tile(np.arange(16, dtype=np.uint32), 4)[None], 4)
    s = np.array([[0.5]], np.float32)
    b = np.array([[-1.0]], np.float32)
    for mode in (0, 1, 2):
        d = ops.dequantize(wq, s, b, 64, 4, "float16", mode)
        assert np.array_equal(d[0, :16], np.arange(16) * 0.5 - 1.0)


def test_dequantize_mode_differences_are_bounded_by_the_product_rounding():
    """The three A.2 candidates differ only by how q*s is rounded before the add: |diff| <= ulp_f16(|q*s|)
    (many result-ulps when q*s + b cancels, which is why the mode must be pinned, not guessed)."""
    rng = np.random.default_rng(1)
    q = rng.integers(0, 16, (64, 256)).astype(np.uint32)
    wq = ops.pack_codes(q, 4)
    s = ops.round_to(rng.uniform(-0.1, 0.1, (64, 4)), "float16")
    b = ops.round_to(rng.uniform(-1, 1, (64, 4)), "float16")
    a = ops.dequantize(wq, s, b, 64, 4, "float16", ops.MODE_FMA)
    c = ops.dequantize(wq, s, b, 64, 4, "float16", ops.MODE_T_MULADD)
    f = ops.dequantize(wq, s, b, 64, 4, "float16", ops.MODE_F32_MULADD)
    prod = np.abs(q.astype(np.float32) * np.repeat(s, 64, -1))
    bound = (prod + np.abs(np.repeat(b, 64, -1))) * 2.0 ** -10 + 1e-7
    assert (np.abs(a - c) <= bound).all()
    assert (a != c).any()                                   # the modes are genuinely distinguishable in f16
    # exact-product-then-f32-add differs from a true fma only through double rounding: rare
    assert (ops.storage_bits(f, "float16") != ops.storage_bits(a, "float16")).mean() < 1e-3


def test_bf16_rounding_half_to_even():
    x = np.array([1.0 + 2 ** -8, 1.0 + 3 * 2 ** -8, 1.0 + 2 ** -8 + 2 ** -20], np.float32)
    r = ops.bf16_round(x)
    assert r[0] == 1.0 and r[1] == np.float32(1.0 + 2 ** -6) and r[2] == np.float32(1.0 + 2 ** -7)


# ---- quantized_matmul (A.3) --------------------------------------------------------------------------
def test_quantized_matmul_matches_grouped_float64():
    rng = np.random.default_rng(2)
    for bits in (4, 8):
        q = rng.integers(0, 1 << bits, (24, 256)).astype(np.uint32)
        wq = ops.pack_codes(q, bits)
        s = rng.uniform(-0.1, 0.1, (24, 4)).astype(np.float32)
        b = rng.uniform(-1, 1, (24, 4)).astype(np.float32)
        x = rng.standard_normal((5, 256)).astype(np.float32)
        y = ops.quantized_matmul(x, wq, s, b, 64, bits)
        y64 = ops.quantized_matmul_grouped(x, wq, s, b, 64, bits)
        assert rel_err(y, y64) < 1e-5


# ---- rms_norm / rope / sdpa (A.4-A.6) ---------------------------------------------------------------
def test_rms_norm_known():
    x = np.array([[3.0, 4.0]], np.float32)          # mean square = 12.5
    y = ops.rms_norm(x, np.array([1.0, 2.0], np.float32), 0.0)
    assert np.allclose(y, [[3 / np.sqrt(12.5), 8 / np.sqrt(12.5)]], rtol=1e-6)


@pytest.mark.parametrize("offset", [0, 1, 255, 8191])
def test_rope_non_traditional_pairing(offset):
    D = 8
    x = np.zeros((1, 1, 1, D), np.float32)
    x[..., 1] = 1.0                                  # only element i=1 set -> pairs with i + D/2 = 5
    y = ops.rope(x, offset, base=10000.0)[0, 0, 0]
    th = offset * 10000.0 ** (-1 / 4)
    exp = np.zeros(D, np.float32)
    exp[1], exp[5] = np.cos(th), np.sin(th)
    assert np.allclose(y, exp, atol=2e-3 if offset > 1000 else 1e-5)
    assert y[0] == 0 and y[4] == 0 and y[2] == 0


def test_rope_per_row_offsets_and_positions():
    rng = np.random.default_rng(3)
    x = rng.standard_normal((2, 3, 4, 16)).astype(np.float32)
    y = ops.rope(x, np.array([5, 9]), 10000.0)
    y0 = ops.rope(x[:1], 5, 10000.0)
    y1 = ops.rope(x[1:], 9, 10000.0)
    assert np.array_equal(y[:1], y0) and np.array_equal(y[1:], y1)
    # position l of a chunk at offset o == single token at offset o + l
    z = ops.rope(x[:1, :, 2:3], 7, 10000.0)
    assert np.allclose(y0[:, :, 2:3], z, atol=1e-6)


def test_sdpa_gqa_head_mapping():
    # 4 q heads over 2 kv heads: q heads 0,1 -> kv 0; 2,3 -> kv 1 (h // (Hq/Hkv))
    B, Hq, Hkv, T, D = 1, 4, 2, 3, 4
    q = np.zeros((B, Hq, 1, D), np.float32)
    k = np.zeros((B, Hkv, T, D), np.float32)
    v = np.zeros((B, Hkv, T, D), np.float32)
    v[0, 0] = 1.0
    v[0, 1] = 2.0
    o = ops.sdpa(q, k, v, 1.0)
    assert np.allclose(o[0, :2], 1.0) and np.allclose(o[0, 2:], 2.0)


def test_sdpa_softmax_known():
    q = np.array([[[[1.0, 0.0]]]], np.float32)
    k = np.array([[[[0.0, 0.0], [np.log(3.0), 0.0]]]], np.float32)
    v = np.array([[[[0.0, 0.0], [4.0, 8.0]]]], np.float32)
    o = ops.sdpa(q, k, v, 1.0)                       # weights 1/4, 3/4
    assert np.allclose(o, [[[[3.0, 6.0]]]], rtol=1e-6)


def test_causal_mask_variable_matches_rule():
    m = ops.additive_causal_mask_variable(3, [2, 0], 5)
    assert m.shape == (2, 3, 5)
    assert (m[0, 0] == [0, 0, 0, -1e9, -1e9]).all()   # offset 2, query 0 sees keys 0..2
    assert (m[1, 0] == [0, -1e9, -1e9, -1e9, -1e9]).all()
    assert (m[1, 2] == [0, 0, 0, -1e9, -1e9]).all()


# ---- sampling (A.7, sample_utils.py) -----------------------------------------------------------------
def test_argmax_lowest_index_on_ties():
    assert ops.argmax_tokens(np.array([[1.0, 3.0, 3.0, 2.0]]))[0, 0] == 1


def test_top_p_distribution_inclusive_prefix_rule():
    p = np.array([[0.5, 0.3, 0.15, 0.05]], np.float32)
    lg = np.log(p)
    d = ops.top_p_distribution(lg, 0.81, 1.0)         # cum = .5, .8, .95, 1 -> keep first two (<= 0.81)
    assert np.allclose(d, [[0.625, 0.375, 0, 0]], atol=1e-6)
    d = ops.top_p_distribution(lg, 0.79, 1.0)         # only the first
    assert np.allclose(d, [[1, 0, 0, 0]], atol=1e-6)
    d = ops.top_p_distribution(lg, 0.4, 1.0)          # empty nucleus -> documented argmax deviation
    assert np.allclose(d, [[1, 0, 0, 0]])


# ---- caches / model graph ----------------------------------------------------------------------------
def test_paged_cache_per_row_offsets_and_growth():
    c = OraclePagedKVCache(4, 2, batch_size=2)
    c.step = 4
    k = np.ones((2, 2, 3, 4), np.float32)
    kf, vf = c.update_and_fetch(k, 2 * k)
    assert kf.shape == (2, 2, 3, 4) and c.keys.shape[2] == 4 and c.offsets == [3, 3]
    kf, vf = c.update_and_fetch(k[:, :, :2], k[:, :, :2])
    assert kf.shape[2] == 5 and c.keys.shape[2] == 8 and c.offsets == [5, 5]
    c.reset()
    assert c.offsets == [0, 0] and c.keys is not None
    c.reset(3)
    assert c.keys is None and c.batch_size == 3


@pytest.mark.parametrize("name,dtype", [("tiny-llama", "float16"), ("tiny-qwen3", "bfloat16")])
def test_prefill_then_decode_equals_full_forward(name, dtype):
    cfg, om, _ = make_pair(name, dtype=dtype, device=None, numerics="mirror")
    ids = prompts_for(cfg, 2, 9)
    full = om.forward(ids, om.make_cache(2))
    cache = om.make_cache(2)
    a = om.forward(ids[:, :6], cache)
    outs = [a]
    for t in range(6, 9):
        outs.append(om.forward(ids[:, t:t + 1], cache))
    inc = np.concatenate(outs, 1)
    assert rel_err(inc, full) < 2e-3 if dtype == "bfloat16" else rel_err(inc, full) < 5e-4


def test_reference_and_mirror_numerics_agree_within_north_star_tolerance():
    cfg, om_ref, _ = make_pair("tiny-llama", dtype="float16", device=None, numerics="reference")
    _, om_mir, _ = make_pair("tiny-llama", dtype="float16", device=None, numerics="mirror")
    ids = prompts_for(cfg, 3, 10)
    a = om_ref.forward(ids, om_ref.make_cache(3), last_only=True)
    b = om_mir.forward(ids, om_mir.make_cache(3), last_only=True)
    assert rel_err(b, a) < 1e-2


def test_generate_step_greedy_is_deterministic_and_advances_cache():
    cfg, om, _ = make_pair("tiny-llama", dtype="float16", device=None)
    p = prompts_for(cfg, 2, 5)
    cache = om.make_cache(2)
    g = generate_step(p, om, cache=cache)
    toks = [next(g)[0] for _ in range(4)]
    assert cache[0].offsets == [8, 8]                 # 5 prompt + 3 fed-back tokens
    g2 = generate_step(p, om)
    toks2 = [next(g2)[0] for _ in range(4)]
    assert all(np.array_equal(a, b) for a, b in zip(toks, toks2))


def test_e4m3_rounding_known_answers():
    """oracle.mlx_ops.e4m3_round (used only by scripts/fp8_kv_error_study.py): OCP E4M3 grid, half-to-even, saturation,
    subnormal step 2^-9."""
    from oracle.mlx_ops import e4m3_round, round_to
    x = np.array([0, 1, 1.0625, 1.125, 1.1875, 448, 500, -0.3, 2.0 ** -9, 2.0 ** -10, 0.0019, 17, 19, 20.1, -1e9], np.float32)
    e = np.array([0, 1, 1.0, 1.125, 1.25, 448, 448, -0.3125, 2.0 ** -9, 0, 2.0 ** -9, 16, 20, 20, -448], np.float32)
    assert np.array_equal(e4m3_round(x), e)
    grid = e4m3_round(np.linspace(-450, 450, 200001).astype(np.float32))
    assert len(np.unique(np.abs(grid))) <= 127 and np.array_equal(e4m3_round(grid), grid)      # idempotent, <= 2^7 - 1 magnitudes
    assert np.array_equal(round_to(x, "float8_e4m3"), e)
