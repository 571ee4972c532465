"""BASELINE-size checks through size-independent properties (the oracle cannot run these shapes in seconds):
agreement between independent kernels of the library, linearity, permutation invariance, idempotence."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _qweights(N, K, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randint(-2 ** 31, 2 ** 31, (N, K // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    s = ((torch.rand((N, K // 64), generator=g, device="cuda") + 0.5) * (2 / 15) / K ** 0.5).to(torch.float16)
    b = (-s.float() * 7.5).to(torch.float16)
    return w, s, b


@pytest.mark.parametrize("N,K,M", [(28672, 4096, 64), (4096, 14336, 64), (128256, 4096, 8), (6144, 4096, 1)])
def test_qmm_paths_agree_and_are_linear_at_llama3_8b_shapes(N, K, M):
    from mlx_parallm_b200 import _lib, ops
    L = _lib.lib()
    w, s, b = _qweights(N, K, N + K)
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn((M, K), generator=g, device="cuda").to(torch.float16)
    ws = torch.zeros((max(ops.qmm_workspace_bytes(M, N, K), 16),), dtype=torch.uint8, device="cuda")
    outs = {}
    for path in (1, 2, 3):
        if path == 3 and M > 16:
            continue
        if path == 1 and M > 8:
            continue                                  # the CUDA-core GEMV at M = 64 takes seconds on these shapes
        _lib.check(L.b200p_qmm_set_path(path))
        try:
            outs[path] = ops.qmm(x, w, s, b, 64, 4, _lib.EPI_STORE_F32, workspace=ws)
        finally:
            L.b200p_qmm_set_path(0)
    ref = outs[2].double()
    scale = ref.abs().max().item()
    for path, o in outs.items():
        assert (o.double() - ref).abs().max().item() <= 2e-3 * scale, path
    # linearity: y(2x) == 2 y(x) (scaling by 2 is exact in f16 and in every fp32 partial sum), y(x1 + x2) ~ y(x1) + y(x2)
    _lib.check(L.b200p_qmm_set_path(2))
    try:
        y2 = ops.qmm((x * 2).contiguous(), w, s, b, 64, 4, _lib.EPI_STORE_F32, workspace=ws)
        x2 = torch.randn((M, K), generator=g, device="cuda").to(torch.float16)
        xs = (x.float() + x2.float()).to(torch.float16)
        ysum = ops.qmm(xs, w, s, b, 64, 4, _lib.EPI_STORE_F32, workspace=ws)
        yb = ops.qmm(x2, w, s, b, 64, 4, _lib.EPI_STORE_F32, workspace=ws)
    finally:
        L.b200p_qmm_set_path(0)
    assert (y2.double() - 2 * ref).abs().max().item() <= 1e-5 * scale
    lin_ref = ops.qmm((xs.float() - x.float()).to(torch.float16), w, s, b, 64, 4, _lib.EPI_STORE_F32, workspace=ws)   # = y(fl16(x1+x2) - x1)
    assert (ysum.double() - ref - lin_ref.double()).abs().max().item() <= 2e-3 * scale
    del yb


@pytest.mark.parametrize("B,T,Hq,Hkv", [(4, 8192, 32, 8), (64, 256, 32, 8)])
def test_decode_attention_variants_agree_and_ignore_key_order(B, T, Hq, Hkv):
    """8-warp split-K, 2-warp and generic CUDA-core kernels agree at an 8k context; softmax attention does not depend on
    the order of the (key, value) pairs."""
    import os
    from mlx_parallm_b200 import _lib, ops
    L = _lib.lib()
    D = 128
    g = torch.Generator(device="cuda").manual_seed(8)
    q = torch.randn((B, 1, Hq, D), generator=g, device="cuda").to(torch.float16)
    k = (torch.randn((B, Hkv, T + 8, D), generator=g, device="cuda") * 0.5).to(torch.float16)
    v = torch.randn((B, Hkv, T + 8, D), generator=g, device="cuda").to(torch.float16)
    lengths = torch.full((B,), T - 1, dtype=torch.int32, device="cuda")       # the kernels see lengths + 1 keys
    lengths[0] = T // 3
    ws = torch.empty((max(ops.attn_workspace_bytes(B, 1, Hq, D), 16),), dtype=torch.uint8, device="cuda")

    def run(path, kk, vv):
        out = torch.zeros((B, 1, Hq * D), dtype=torch.float16, device="cuda")
        _lib.check(L.b200p_attn_set_path(path))
        try:
            ops.attn(q, kk, vv, lengths, out, B, 1, Hq, Hkv, D, D ** -0.5, workspace=ws)
        finally:
            L.b200p_attn_set_path(0)
        return out.float()

    a = run(2, k, v)
    gen = run(1, k, v)
    assert (a - gen).abs().max().item() <= 2e-3 * gen.abs().max().item()
    perm = torch.randperm(T, generator=torch.Generator().manual_seed(1)).cuda()
    kp, vp = k.clone(), v.clone()
    kp[1:, :, :T] = k[1:, :, perm]                    # rows 1.. hold T valid keys; row 0 (shorter) is left alone
    vp[1:, :, :T] = v[1:, :, perm]
    ap = run(2, kp, vp)
    assert (ap - a).abs().max().item() <= 2e-3 * a.abs().max().item()


def test_prefill_attention_matches_decode_attention_row_by_row():
    """The last query of a causal prefill chunk sees exactly the keys a decode step at that position sees."""
    from mlx_parallm_b200 import _lib, ops
    B, Lq, Hq, Hkv, D, T0 = 2, 96, 32, 8, 128, 1000
    g = torch.Generator(device="cuda").manual_seed(12)
    q = torch.randn((B, Lq, Hq, D), generator=g, device="cuda").to(torch.float16)
    k = (torch.randn((B, Hkv, T0 + Lq + 8, D), generator=g, device="cuda") * 0.5).to(torch.float16)
    v = torch.randn((B, Hkv, T0 + Lq + 8, D), generator=g, device="cuda").to(torch.float16)
    lengths = torch.tensor([T0, T0 - 333], dtype=torch.int32, device="cuda")
    ws = torch.empty((max(ops.attn_workspace_bytes(B, 1, Hq, D), 16),), dtype=torch.uint8, device="cuda")
    pre = torch.zeros((B, Lq, Hq * D), dtype=torch.float16, device="cuda")
    ops.attn(q, k, v, lengths, pre, B, Lq, Hq, Hkv, D, D ** -0.5, workspace=ws)
    for l in (0, 63, 64, Lq - 1):
        dec = torch.zeros((B, 1, Hq * D), dtype=torch.float16, device="cuda")
        ops.attn(q[:, l:l + 1].contiguous(), k, v, lengths + l, dec, B, 1, Hq, Hkv, D, D ** -0.5, workspace=ws)
        assert (dec[:, 0].float() - pre[:, l].float()).abs().max().item() <= 2e-3 * dec.float().abs().max().item(), l
