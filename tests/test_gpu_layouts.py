"""PERM8 activation layout and the fp32-accumulating fused projections: every producer / consumer of the permuted
layout against the natural-order oracle, and the self-zeroing consumers (rope_kv_append on a float buffer, silu_mul)."""
import numpy as np
import pytest
import torch

from oracle import mlx_ops as ref
from helpers import rel_err
from test_gpu_kernels import TD, attn_case, t_, w_

pytestmark = pytest.mark.gpu

PERM = np.array([0, 4, 1, 5, 2, 6, 3, 7])


def to_perm8(a):
    """natural -> PERM8 along the last axis: stored[8j + p] = natural[8j + PERM[p]]."""
    s = a.shape
    return a.reshape(*s[:-1], s[-1] // 8, 8)[..., PERM].reshape(s)


def from_perm8(a):
    s = a.shape
    out = np.empty_like(a).reshape(*s[:-1], s[-1] // 8, 8)
    out[..., PERM] = a.reshape(*s[:-1], s[-1] // 8, 8)
    return out.reshape(s)


@pytest.fixture(scope="module")
def ops():
    from mlx_parallm_b200 import ops as o, _lib
    _lib.require_device()
    return o


def test_perm8_helpers_roundtrip():
    a = np.arange(32.0).reshape(2, 16)
    assert np.array_equal(from_perm8(to_perm8(a)), a)
    assert list(to_perm8(np.arange(8.0))) == [0, 4, 1, 5, 2, 6, 3, 7]


@pytest.mark.parametrize("dtype", ["float16", "bfloat16"])
def test_rmsnorm_perm8(ops, dtype):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((5, 512)).astype(np.float32) * 3
    w = ref.round_to(1 + 0.1 * rng.standard_normal(512), dtype)
    nat = ops.rmsnorm(torch.from_numpy(x).cuda(), t_(w, dtype), 1e-5).float().cpu().numpy()
    per = ops.rmsnorm(torch.from_numpy(x).cuda(), t_(w, dtype), 1e-5, out_perm=True).float().cpu().numpy()
    assert np.array_equal(from_perm8(per), nat)          # same values, permuted columns: bit-exact


@pytest.mark.parametrize("path,L", [(1, 1), (2, 1), (2, 70)])
def test_attention_perm8_output(ops, path, L):
    from mlx_parallm_b200 import _lib
    B, Hq, Hkv, D, lengths = 3, 8, 2, 128, [0, 50, 700]
    nat, _ = attn_case(ops, B, L, Hq, Hkv, D, lengths, "float16", path=path, seed=2)
    rng = np.random.default_rng(2)
    Tcap = int(max(lengths)) + L + 3
    q = ref.round_to(rng.standard_normal((B, L, Hq, D)), "float16")
    k = ref.round_to(rng.standard_normal((B, Hkv, Tcap, D)), "float16")
    v = ref.round_to(rng.standard_normal((B, Hkv, Tcap, D)), "float16")
    out = torch.zeros((B, L, Hq * D), dtype=torch.float16, device="cuda")
    ws = torch.empty((max(ops.attn_workspace_bytes(B, L, Hq, D), 16),), dtype=torch.uint8, device="cuda")
    _lib.check(_lib.lib().b200p_attn_set_path(path))
    try:
        ops.attn(t_(q, "float16"), t_(k, "float16"), t_(v, "float16"), torch.tensor(lengths, dtype=torch.int32).cuda(), out,
                 B, L, Hq, Hkv, D, D ** -0.5, workspace=ws, out_perm=True)
    finally:
        _lib.lib().b200p_attn_set_path(0)
    assert np.array_equal(from_perm8(out.float().cpu().numpy()), nat)


def _qmm_inputs(M, N, K, dtype, seed):
    rng = np.random.default_rng(seed)
    wq = rng.integers(0, 2 ** 32, (N, K // 8), dtype=np.uint32)
    s = ref.round_to(rng.uniform(0.5, 1.5, (N, K // 64)) * rng.choice([-1, 1], (N, K // 64)) * (2 / 15) / np.sqrt(K), dtype)
    b = ref.round_to(-s * 7.5 + rng.normal(0, 0.01 / np.sqrt(K), (N, K // 64)), dtype)
    x = ref.round_to(rng.standard_normal((M, K)), dtype)
    return wq, s, b, x


@pytest.mark.parametrize("M,N,K,dtype,path", [
    (64, 1024, 512, "float16", 2), (16, 512, 1024, "bfloat16", 2), (64, 256, 4096, "float16", 2),   # tensor path
    (3, 128, 256, "float16", 1), (8, 96, 512, "bfloat16", 1),                                       # GEMV path
    (3, 128, 256, "float16", 3), (12, 512, 1024, "bfloat16", 3), (1, 64, 4096, "float16", 3),       # mma.sync GEMV
])
def test_qmm_x_perm8_equals_natural(ops, M, N, K, dtype, path):
    from mlx_parallm_b200 import _lib
    wq, s, b, x = _qmm_inputs(M, N, K, dtype, seed=M + N)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    _lib.check(_lib.lib().b200p_qmm_set_path(path))
    try:
        nat = ops.qmm(t_(x, dtype), w_(wq), t_(s, dtype), t_(b, dtype), 64, 4, _lib.EPI_STORE_F32, workspace=ws)
        per = ops.qmm(t_(to_perm8(x), dtype), w_(wq), t_(s, dtype), t_(b, dtype), 64, 4, _lib.EPI_STORE_F32, workspace=ws,
                      x_perm=True)
    finally:
        _lib.lib().b200p_qmm_set_path(0)
    exp = ref.quantized_matmul(x, wq, s, b, 64, 4)
    tol = {"float16": 3e-3, "bfloat16": 2e-2}[dtype]
    assert rel_err(per.cpu().numpy(), exp) < tol
    assert rel_err(per.cpu().numpy(), nat.cpu().numpy()) < 1e-5       # same products, different summation order


@pytest.mark.parametrize("M,N,K,epi,path", [(64, 2048, 1024, 3, 2), (64, 1024, 512, 0, 2), (40, 512, 4096, 0, 2), (4, 256, 256, 3, 1)])
def test_qmm_y_perm8_output(ops, M, N, K, epi, path):
    from mlx_parallm_b200 import _lib
    wq, s, b, x = _qmm_inputs(M, N, K, "float16", seed=N + K)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    _lib.check(_lib.lib().b200p_qmm_set_path(path))
    try:
        nat = ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, epi, workspace=ws)
        per = ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, epi, workspace=ws, y_perm=True)
    finally:
        _lib.lib().b200p_qmm_set_path(0)
    # split tiles are summed in arrival order, so two launches agree to fp32 rounding before the final 16-bit rounding
    assert rel_err(from_perm8(per.float().cpu().numpy()), nat.float().cpu().numpy()) < 1.1e-3


def test_qmm_accumulates_into_zeroed_f32_and_silu_mul_rezeroes(ops):
    from mlx_parallm_b200 import _lib
    M, I, K = 64, 1024, 1024
    wq, s, b, x = _qmm_inputs(M, 2 * I, K, "float16", seed=5)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, 2 * I, K), 16),), dtype=torch.uint8, device="cuda")
    gu = torch.zeros((M, 2 * I), dtype=torch.float32, device="cuda")
    act = torch.empty((M, I), dtype=torch.float16, device="cuda")
    acc = ref.quantized_matmul(x, wq, s, b, 64, 4).astype(np.float64)
    exp = acc[:, :I] / (1 + np.exp(-acc[:, :I])) * acc[:, I:]
    for perm in (False, True, False):                  # the buffer must come back zero every time
        ops.qmm(t_(x, "float16"), w_(wq), t_(s, "float16"), t_(b, "float16"), 64, 4, _lib.EPI_ADD_RESIDUAL, out=gu, workspace=ws)
        ops.silu_mul(gu, act, out_perm=perm)
        got = act.float().cpu().numpy()
        assert rel_err(from_perm8(got) if perm else got, exp) < 3e-3
        assert not gu.any().item()


@pytest.mark.parametrize("dtype,D,qk_norm", [("float16", 128, False), ("bfloat16", 128, True)])
def test_rope_kv_append_from_f32_buffer(ops, dtype, D, qk_norm):
    rng = np.random.default_rng(4)
    B, L, Hq, Hkv, Tcap = 3, 2, 4, 2, 40
    lengths = torch.tensor([0, 7, 33], dtype=torch.int32).cuda()
    qkv = ref.round_to(rng.standard_normal((B * L, (Hq + 2 * Hkv) * D)), dtype)
    qn, kn = (t_(ref.round_to(1 + 0.1 * rng.standard_normal(D), dtype), dtype) for _ in range(2))

    def run(src):
        q_out = torch.zeros((B, L, Hq, D), dtype=TD[dtype], device="cuda")
        kc = torch.zeros((B, Hkv, Tcap, D), dtype=TD[dtype], device="cuda")
        vc = torch.zeros_like(kc)
        ops.rope_kv_append(src, q_out, kc, vc, lengths, B, L, Hq, Hkv, D, 10000.0, 1.0,
                           qn if qk_norm else None, kn if qk_norm else None, 1e-6)
        return q_out, kc, vc

    a = run(t_(qkv, dtype))
    # float input carrying sub-ulp noise: it must be rounded to the storage type first, then behave identically
    noisy = torch.from_numpy((qkv * (1 + 1e-5)).astype(np.float32)).cuda()
    b = run(noisy)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    assert not noisy.any().item()                      # re-zeroed behind the read
